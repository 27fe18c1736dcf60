mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -2
PADNE_B200_TRACE=1 timeout 600 python profiles/precond_sweep.py 2layer_1M full 1600 2>&1 | grep -v jacobi | tail -2 | cut -c1-400
timeout 600 python bench.py --steps 5 --warmup 3 --no-cpu-baseline 2>&1 | python -c "
import sys,json
for l in sys.stdin:
    if not l.startswith('{'): continue
    d=json.loads(l); r=d['roofline']; print('value',round(d['value']/1e6,2),'ms/step',round(d['ms_per_step'],2),'e2e',round(d['e2e']['value']/1e6,2),'frac',round(r['frac'],3),'insitu',r['ms_per_launch_in_situ'],'warm',r['ms_per_launch_warm_l2'],'cold',r['ms_per_launch_cold_l2'], 'it', d['config']['iterations'])"
