# final-tree check: GPU suite, smoke, default bench line
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -x -q 2>&1 | tail -3
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -2
timeout 900 python bench.py > gpurun_out/r2_bench_final.log 2>&1; echo "bench exit $?"
grep '^{' gpurun_out/r2_bench_final.log | python -c "
import sys,json
d=json.loads(sys.stdin.read()); c=d['config']; print(round(d['value']/1e6,2),'MDOF/s',round(d['ms_per_step'],2),'ms iters',c['iterations'],'t_coarse',round(c['t_coarse_ms'],3),'setup',round(c['t_setup_ms'],2),'us/it',round(c['us_per_iteration'],1),'e2e',round(d['e2e']['value']/1e6,2),'frac',round(d['roofline']['frac'],3))
for k,v in d.get('named_configs',{}).items(): print('   ',k, v['n_gpus'],'GPUs', round(v['ms_per_step'],2),'ms', v['iterations'],'iters', 't_coarse', round(v['t_coarse_ms'],2), v['method'], 'res', v['residual_norm'])"
