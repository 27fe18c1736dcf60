# multi-GPU check of a changed tree: partition parity with and without the rank-local level, then the default bench line
N=${1:-2}
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29533"
for ll in 0 1; do
PADNE_B200_LOCAL_LEVEL=$ll timeout 600 $TR scripts/dist_check.py 2layer_1M 300 8layer_10M 300 4layer_regulator_4M 120 > gpurun_out/dist_check_${N}_ll$ll.log 2>&1; echo "dist_check LOCAL_LEVEL=$ll exit $?"; grep "^{" gpurun_out/dist_check_${N}_ll$ll.log | python -c "
import sys,json
for l in sys.stdin:
    d=json.loads(l); print(d['case'],d['grid'],d['method'],'kind',d['kind_multi'],'it',d['iters_single'],d['iters_multi'],'us/it',d['us_per_iter_multi'],'ncl',d['n_coarse_local'],'diff',d['max_rel_diff_vs_single'],'res',d['residual_norm_multi'],'ok',d['ok'])"
grep -iE "error|Traceback" gpurun_out/dist_check_${N}_ll$ll.log | head -5
done
bash scripts/gpu_r2_scale.sh $N
