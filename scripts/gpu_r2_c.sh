# round 2, third GPU pass: v2 with 38-bit restriction on both SpMV kernels, against v1
mkdir -p gpurun_out
timeout 300 python __graft_entry__.py --smoke 2>&1 | tail -1
timeout 600 python scripts/gpu_diag.py v2ring > gpurun_out/diag_v2ring.jsonl 2> gpurun_out/diag_v2ring.err; echo "diag v2ring exit $?"
PADNE_B200_SPMV=xs timeout 600 python scripts/gpu_diag.py v2xs > gpurun_out/diag_v2xs.jsonl 2> gpurun_out/diag_v2xs.err; echo "diag v2xs exit $?"
timeout 900 python -m pytest tests -m gpu -q 2>&1 | tail -40 > gpurun_out/pytest_c.log; tail -3 gpurun_out/pytest_c.log
timeout 600 python bench.py --steps 5 --warmup 3 > gpurun_out/bench_v2ring.log 2>&1; echo "bench v2ring exit $?"
PADNE_B200_SPMV=xs timeout 600 python bench.py --steps 5 --warmup 3 --no-cpu-baseline --no-named-configs > gpurun_out/bench_v2xs.log 2>&1; echo "bench v2xs exit $?"
PADNE_B200_PCG=v1 timeout 600 python bench.py --steps 5 --warmup 3 --no-cpu-baseline --no-named-configs > gpurun_out/bench_v1ring.log 2>&1; echo "bench v1ring exit $?"
for f in gpurun_out/bench_v2ring.log gpurun_out/bench_v2xs.log gpurun_out/bench_v1ring.log; do grep '^{' $f | python -c "
import sys,json
for l in sys.stdin:
    d=json.loads(l); c=d['config']; r=d['roofline']
    print(d['value']/1e6,'MDOF/s', d['ms_per_step'],'ms', 'iters',c['iterations'],'us/it',round(c['us_per_iteration'],1),'res',c['residual_norm'],'spmv us',round(1e3*r['ms_per_launch_in_situ'],2),'frac',round(r['frac'],3),'e2e',d['e2e']['value']/1e6, 'setup',c['t_setup_ms'],'coarse',c['t_coarse_ms'])
    for k,v in d.get('named_configs',{}).items(): print('  ',k, v['ms_per_step'],'ms', v['iterations'],'iters', v['method'], 'res', v['residual_norm'])
"; done
PADNE_B200_KTIME=1 timeout 300 python profiles/one_solve.py 2>&1 | grep -A12 ktime | head -16
