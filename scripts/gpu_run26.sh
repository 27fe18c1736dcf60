mkdir -p gpurun_out
N=${1:-2}
PADNE_B200_PEER_STATS=1 timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus $N --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/bench$N.log 2>&1; echo "bench exit $?" >> gpurun_out/bench$N.log
grep -E "exit|exchanges|intervals" gpurun_out/bench$N.log | tail -6; grep '^{' gpurun_out/bench$N.log | python -c "
import sys,json
for l in sys.stdin:
    d=json.loads(l); c=d['config']; print(d['n_gpus'], c['workload'], d['scaling'], 'value', round(d['value']/1e6,2), 'MDOF/s ms/step', round(d['ms_per_step'],2), 'it', c['iterations'], 'krylov', round(c['t_krylov_ms'],2), 'us/it', round(1e3*c['t_krylov_ms']/c['iterations'],1), 'setup', round(c['t_setup_ms'],1), 'e2e', round(d['e2e']['value']/1e6,2), 'frac', round(d['roofline']['frac'],3), 'launches', d['gpu_launches'])
"
