# multi-GPU pass (N = $1): parity against the single-GPU solve, then the weak-scaling bench line for v2 and v1
N=${1:-2}
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29533"
timeout 600 $TR scripts/dist_check.py 2layer_1M 300 8layer_10M 300 4layer_regulator_4M 120 2layer_1M 708 > gpurun_out/dist_check_$N.log 2>&1; echo "dist_check exit $?"; grep "^{" gpurun_out/dist_check_$N.log | cut -c1-700; grep -iE "error|Traceback" gpurun_out/dist_check_$N.log | head -5
for pcg in v2 v1; do
PADNE_B200_PCG=$pcg PADNE_B200_PEER_STATS=1 timeout 600 $TR bench.py --gpus $N --steps 5 --warmup 3 --no-cpu-baseline --no-named-configs > gpurun_out/bench_${N}_$pcg.log 2>&1; echo "bench $pcg exit $?"
grep "event intervals" gpurun_out/bench_${N}_$pcg.log | head -1
grep '^{' gpurun_out/bench_${N}_$pcg.log | python -c "
import sys,json
for l in sys.stdin:
    d=json.loads(l); c=d['config']
    print('$pcg', d['value']/1e6,'MDOF/s', round(d['ms_per_step'],2),'ms', 'iters',c['iterations'],'us/it',round(c['us_per_iteration'],1),'kind',c['iteration_kind'],'nc',c['n_coarse'],c['n_coarse_local'],c['local_level'],'res',c['residual_norm'],'e2e',d['e2e']['value']/1e6, 'setup',round(c['t_setup_ms'],2),'coarse',round(c['t_coarse_ms'],2))
"
done
tail -5 gpurun_out/bench_${N}_v2.log | cut -c1-300
