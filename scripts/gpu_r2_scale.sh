# the driver's scaling command at N ranks (default workload + named configs), builder-run
N=${1:-2}
mkdir -p gpurun_out
if [ "$N" = "1" ]; then TR="python"; else TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29533"; fi
( time timeout 1200 $TR bench.py --gpus $N --steps 10 --warmup 3 > gpurun_out/r2_scale_$N.log 2>&1 ) 2>&1 | grep real; echo "bench exit $?"
grep -E "Error|error|Traceback" gpurun_out/r2_scale_$N.log | head -5
grep '^{' gpurun_out/r2_scale_$N.log | python -c "
import sys,json
for l in sys.stdin:
    d=json.loads(l); c=d['config']
    print(d['n_gpus'],'GPUs', round(d['value']/1e6,2),'MDOF/s', round(d['ms_per_step'],2),'ms', 'iters',c['iterations'],'us/it',round(c['us_per_iteration'],1),'kind',c['iteration_kind'],'nc',c['n_coarse'],c['n_coarse_local'],c['local_level'],'res',c['residual_norm'],'e2e',round(d['e2e']['value']/1e6,2), 'setup',round(c['t_setup_ms'],2),'krylov',round(c['t_krylov_ms'],2))
    for k,v in d.get('named_configs',{}).items(): print('   ',k, v['n_gpus'],'GPUs', round(v['ms_per_step'],2),'ms', v['iterations'],'iters', round(v['us_per_iteration'],1),'us/it', v['method'], 'res', v['residual_norm'], round(v['value']/1e6,2),'MDOF/s')
"
