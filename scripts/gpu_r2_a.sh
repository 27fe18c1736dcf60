# round 2, first GPU pass: diagnostics of the new iteration (v2) against the round-1 one (v1) and the round-1 SpMV (ring)
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm --format=csv,noheader
timeout 300 python __graft_entry__.py --smoke 2>&1 | tail -2
timeout 600 python scripts/gpu_diag.py v2 > gpurun_out/diag_v2.jsonl 2> gpurun_out/diag_v2.err; echo "diag v2 exit $?"
PADNE_B200_PCG=v1 timeout 600 python scripts/gpu_diag.py v1xs > gpurun_out/diag_v1xs.jsonl 2> gpurun_out/diag_v1xs.err; echo "diag v1xs exit $?"
PADNE_B200_SPMV=ring timeout 600 python scripts/gpu_diag.py v1ring > gpurun_out/diag_v1ring.jsonl 2> gpurun_out/diag_v1ring.err; echo "diag v1ring exit $?"
timeout 900 python -m pytest tests -m gpu -q 2>&1 | tail -40 > gpurun_out/pytest_a.log; tail -5 gpurun_out/pytest_a.log
timeout 600 python bench.py --steps 5 --warmup 3 > gpurun_out/bench_v2.log 2>&1; echo "bench v2 exit $?"
PADNE_B200_PCG=v1 timeout 600 python bench.py --steps 5 --warmup 3 --no-cpu-baseline --no-named-configs > gpurun_out/bench_v1xs.log 2>&1; echo "bench v1xs exit $?"
PADNE_B200_SPMV=ring timeout 600 python bench.py --steps 5 --warmup 3 --no-cpu-baseline --no-named-configs > gpurun_out/bench_v1ring.log 2>&1; echo "bench v1ring exit $?"
for f in gpurun_out/bench_v2.log gpurun_out/bench_v1xs.log gpurun_out/bench_v1ring.log; do grep '^{' $f | python -c "
import sys,json
for l in sys.stdin:
    d=json.loads(l); c=d['config']; r=d['roofline']
    print(d['value']/1e6,'MDOF/s', d['ms_per_step'],'ms', 'iters',c['iterations'],'us/it',round(c['us_per_iteration'],1),'res',c['residual_norm'],'spmv us',round(1e3*r['ms_per_launch_in_situ'],2),'frac',round(r['frac'],3),'e2e',d['e2e']['value']/1e6, 'setup',c['t_setup_ms'],'coarse',c['t_coarse_ms'])
    for k,v in d.get('named_configs',{}).items(): print('  ',k, v['ms_per_step'],'ms', v['iterations'],'iters', v['method'], 'res', v['residual_norm'])
"; done
