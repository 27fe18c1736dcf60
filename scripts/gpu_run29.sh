timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -2
timeout 600 python profiles/precond_sweep.py 2layer_1M full 1600 2>&1 | python -c "
import sys,json
for l in sys.stdin:
    try: d=json.loads(l)
    except Exception: print(l[:200]); continue
    print({k:d.get(k) for k in ('precond','n_coarse','iterations','t_krylov_ms','t_coarse_ms','t_total_ms','us_per_iteration','error')})"
timeout 600 python bench.py --steps 5 --warmup 3 --no-cpu-baseline 2>&1 | python -c "
import sys,json
for l in sys.stdin:
    if not l.startswith('{'): continue
    d=json.loads(l); r=d['roofline']; print('value',round(d['value']/1e6,2),'ms/step',round(d['ms_per_step'],2),'e2e',round(d['e2e']['value']/1e6,2),'frac',round(r['frac'],3),'insitu',r['ms_per_launch_in_situ'],'krylov',d['config']['t_krylov_ms'], 'it', d['config']['iterations'])"
