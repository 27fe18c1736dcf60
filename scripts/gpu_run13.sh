# phase trace of the partitioned solve (N ranks)
mkdir -p gpurun_out
N=${1:-4}
for cfg in "" "--config 8layer_10M"; do
PADNE_B200_TRACE=1 PADNE_B200_PEER_STATS=1 timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus $N --steps 2 --warmup 3 --no-cpu-baseline $cfg > gpurun_out/trace$N.log 2>&1; echo "bench exit $?" >> gpurun_out/trace$N.log
grep -E "trace rank 0|exit" gpurun_out/trace$N.log | tail -4; grep '^{' gpurun_out/trace$N.log | python -c "
import sys,json
for l in sys.stdin:
    d=json.loads(l); c=d['config']; print(d['n_gpus'], c['workload'], d['scaling'], 'value', round(d['value']/1e6,2), 'MDOF/s ms/step', round(d['ms_per_step'],2), 'it', c['iterations'], 'krylov', round(c['t_krylov_ms'],2), 'us/it', round(1e3*c['t_krylov_ms']/c['iterations'],1), 'setup', round(c['t_setup_ms'],1), 'e2e', round(d['e2e']['value']/1e6,2), 'frac', round(d['roofline']['frac'],3))
"
cp gpurun_out/trace$N.log "gpurun_out/trace${N}_$(echo $cfg | tr -d ' -')x.log"
done
