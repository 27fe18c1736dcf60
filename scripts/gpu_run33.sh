mkdir -p gpurun_out
N=${1:-2}
for ll in 0 1; do
echo "== LOCAL_LEVEL=$ll"
PADNE_B200_LOCAL_LEVEL=$ll timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 scripts/dist_check.py 2layer_1M 300 8layer_10M 300 > gpurun_out/dist$N.log 2>&1; echo "dist exit $?" >> gpurun_out/dist$N.log
grep -E '^\{|exit|rror' gpurun_out/dist$N.log | cut -c1-420
for cfg in "" "--config 8layer_10M"; do
PADNE_B200_LOCAL_LEVEL=$ll PADNE_B200_PEER_STATS=1 timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus $N --steps 3 --warmup 3 --no-cpu-baseline $cfg > gpurun_out/bench$N.log 2>&1; echo "bench exit $?" >> gpurun_out/bench$N.log
grep -E "exit|rank 0\] event" gpurun_out/bench$N.log | tail -2; grep '^{' gpurun_out/bench$N.log | python -c "
import sys,json
for l in sys.stdin:
    d=json.loads(l); c=d['config']; print(d['n_gpus'], c['workload'], d['scaling'], 'value', round(d['value']/1e6,2), 'MDOF/s ms/step', round(d['ms_per_step'],2), 'it', c['iterations'], 'krylov', round(c['t_krylov_ms'],2), 'us/it', round(1e3*c['t_krylov_ms']/c['iterations'],1), 'setup', round(c['t_setup_ms'],1), 'coarse', round(c['t_coarse_ms'],2), 'e2e', round(d['e2e']['value']/1e6,2))
"
done
done
