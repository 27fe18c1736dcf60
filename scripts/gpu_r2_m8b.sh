N=${1:-8}
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29533"
timeout 600 $TR scripts/dist_check.py 2layer_1M 300 8layer_10M 300 4layer_regulator_4M 120 > gpurun_out/dist_check_$N.log 2>&1; echo "dist_check exit $?"; grep "^{" gpurun_out/dist_check_$N.log | python -c "
import sys,json
for l in sys.stdin:
    d=json.loads(l); print(d['case'],d['grid'],d['method'],'kind',d['kind_multi'],'it',d['iters_single'],d['iters_multi'],'us/it',d['us_per_iter_multi'],'ncl',d['n_coarse_local'],'diff',d['max_rel_diff_vs_single'],'res',d['residual_norm_multi'],'ok',d['ok'])"
bash scripts/gpu_r2_scale.sh $N
PADNE_B200_PEER_STATS=1 timeout 600 $TR bench.py --gpus $N --steps 3 --warmup 3 --no-cpu-baseline --no-named-configs 2>&1 | grep "event intervals" | head -2
