N=${1:-2}
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29533"
for v in "PADNE_B200_PCG=v1"; do
env $v PADNE_B200_PEER_STATS=1 timeout 600 $TR bench.py --gpus $N --steps 5 --warmup 3 --no-cpu-baseline --no-named-configs > gpurun_out/b.log 2>&1
grep "event intervals" gpurun_out/b.log | head -1
grep '^{' gpurun_out/b.log | python -c "
import sys,json
for l in sys.stdin:
    d=json.loads(l); c=d['config']
    print('$v', round(d['value']/1e6,2),'MDOF/s', round(d['ms_per_step'],2),'ms', 'iters',c['iterations'],'us/it',round(c['us_per_iteration'],1))
"
done
