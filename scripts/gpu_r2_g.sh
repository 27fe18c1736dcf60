mkdir -p gpurun_out
timeout 300 python scripts/spmv_ab.py 2>&1 | tail -1
timeout 900 python -m pytest tests -m gpu -q -x -k "not large_sub and not full_size" 2>&1 | tail -3
timeout 600 python bench.py --steps 5 --warmup 3 --no-cpu-baseline --no-named-configs > gpurun_out/bench_v1.log 2>&1; echo "bench exit $?"
grep '^{' gpurun_out/bench_v1.log | python -c "
import sys,json
for l in sys.stdin:
    d=json.loads(l); c=d['config']; r=d['roofline']
    print(d['value']/1e6,'MDOF/s', d['ms_per_step'],'ms', 'iters',c['iterations'],'us/it',round(c['us_per_iteration'],1),'res',c['residual_norm'],'spmv us',round(1e3*r['ms_per_launch_in_situ'],2),'frac',round(r['frac'],3),'e2e',d['e2e']['value']/1e6, 'setup',c['t_setup_ms'],'coarse',c['t_coarse_ms'])
"
