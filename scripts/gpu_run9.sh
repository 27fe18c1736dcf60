mkdir -p gpurun_out
PADNE_B200_PEER_STATS=1 timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus 2 --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/bench2.log 2>&1; echo "bench2 exit $?" >> gpurun_out/bench2.log
grep -E "padne_b200 rank|exit" gpurun_out/bench2.log; grep '^{' gpurun_out/bench2.log | python -c "
import sys,json
for l in sys.stdin:
    d=json.loads(l); c=d['config']; print(d['n_gpus'], c['workload'], 'value', round(d['value']/1e6,2), 'ms', round(d['ms_per_step'],2), 'it', c['iterations'], 'krylov', round(c['t_krylov_ms'],2), 'us/it', round(1e3*c['t_krylov_ms']/c['iterations'],1), 'launches', d['gpu_launches'])
"
