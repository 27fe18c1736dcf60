# round 2, second GPU pass: v2 after the zmax fix, refinement-style restarts, ncu captures of both SpMV kernels
mkdir -p gpurun_out
timeout 300 python __graft_entry__.py --smoke 2>&1 | tail -1
timeout 600 python scripts/gpu_diag.py v2 > gpurun_out/diag_v2.jsonl 2> gpurun_out/diag_v2.err; echo "diag v2 exit $?"
PADNE_B200_PCG=v1 timeout 600 python scripts/gpu_diag.py v1xs > gpurun_out/diag_v1xs.jsonl 2> gpurun_out/diag_v1xs.err; echo "diag v1xs exit $?"
timeout 900 python -m pytest tests -m gpu -q 2>&1 | tail -40 > gpurun_out/pytest_b.log; tail -3 gpurun_out/pytest_b.log
timeout 600 python bench.py --steps 5 --warmup 3 > gpurun_out/bench_v2.log 2>&1; echo "bench v2 exit $?"
PADNE_B200_SPMV=ring timeout 600 python bench.py --steps 5 --warmup 3 --no-cpu-baseline --no-named-configs > gpurun_out/bench_v1ring.log 2>&1; echo "bench v1ring exit $?"
for f in gpurun_out/bench_v2.log gpurun_out/bench_v1ring.log; do grep '^{' $f | python -c "
import sys,json
for l in sys.stdin:
    d=json.loads(l); c=d['config']; r=d['roofline']
    print(d['value']/1e6,'MDOF/s', d['ms_per_step'],'ms', 'iters',c['iterations'],'us/it',round(c['us_per_iteration'],1),'res',c['residual_norm'],'spmv us',round(1e3*r['ms_per_launch_in_situ'],2),'frac',round(r['frac'],3),'e2e',d['e2e']['value']/1e6, 'setup',c['t_setup_ms'],'coarse',c['t_coarse_ms'])
    for k,v in d.get('named_configs',{}).items(): print('  ',k, v['ms_per_step'],'ms', v['iterations'],'iters', v['method'], 'res', v['residual_norm'])
"; done
ONE_SOLVE_ITERS=64 timeout 600 ncu --set full --clock-control none --import-source on -k regex:sell_xs_kernel -s 40 -c 1 -o gpurun_out/r2_xs -f python profiles/one_solve.py > gpurun_out/ncu_xs.log 2>&1; tail -1 gpurun_out/ncu_xs.log
PADNE_B200_SPMV=ring ONE_SOLVE_ITERS=64 timeout 600 ncu --set full --clock-control none --import-source on -k regex:sell_tma_kernel -s 40 -c 1 -o gpurun_out/r2_ring -f python profiles/one_solve.py > gpurun_out/ncu_ring.log 2>&1; tail -1 gpurun_out/ncu_ring.log
ONE_SOLVE_ITERS=64 timeout 600 ncu --set full --clock-control none --import-source on -k regex:cg_update2 -s 40 -c 1 -o gpurun_out/r2_upd2 -f python profiles/one_solve.py > gpurun_out/ncu_upd2.log 2>&1; tail -1 gpurun_out/ncu_upd2.log
ls -la gpurun_out/*.ncu-rep
