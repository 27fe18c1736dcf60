N=${1:-2}
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29533"
PADNE_B200_PEER_STATS=1 timeout 600 $TR bench.py --gpus $N --steps 5 --warmup 3 --no-cpu-baseline --no-named-configs > gpurun_out/bench_${N}_v2.log 2>&1; echo "bench exit $?"
grep "event intervals" gpurun_out/bench_${N}_v2.log | head -2
grep '^{' gpurun_out/bench_${N}_v2.log | python -c "
import sys,json
for l in sys.stdin:
    d=json.loads(l); c=d['config']
    print(d['value']/1e6,'MDOF/s', round(d['ms_per_step'],2),'ms', 'iters',c['iterations'],'us/it',round(c['us_per_iteration'],1),'nc',c['n_coarse'],c['n_coarse_local'],'res',c['residual_norm'],'e2e',d['e2e']['value']/1e6, 'setup',round(c['t_setup_ms'],2),'coarse',round(c['t_coarse_ms'],2),'krylov',round(c['t_krylov_ms'],2))
"
