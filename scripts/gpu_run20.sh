mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -3
timeout 600 python profiles/precond_sweep.py 2layer_1M full 1600 2>&1 | python -c "
import sys,json
for l in sys.stdin:
    try: d=json.loads(l)
    except Exception: print(l[:200]); continue
    print({k:d.get(k) for k in ('precond','n_coarse','iterations','t_krylov_ms','t_coarse_ms','t_total_ms','us_per_iteration','error')})"
timeout 600 python profiles/precond_sweep.py 8layer_10M full 1600 2>&1 | python -c "
import sys,json
for l in sys.stdin:
    try: d=json.loads(l)
    except Exception: print(l[:200]); continue
    print({k:d.get(k) for k in ('precond','n_coarse','iterations','t_krylov_ms','t_coarse_ms','t_total_ms','us_per_iteration','error')})"
