mkdir -p gpurun_out
timeout 60 python -m pytest tests -m gpu -x -q 2>&1 | tail -2
timeout 40 python bench.py --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/bench_last.log 2>&1; echo "bench exit $?"
grep '^{' gpurun_out/bench_last.log | python -c "
import sys,json
for l in sys.stdin:
    d=json.loads(l); r=d['roofline']; print('value',round(d['value']/1e6,2),'ms/step',round(d['ms_per_step'],2),'e2e',round(d['e2e']['value']/1e6,2),'frac',round(r['frac'],3),'it',d['config']['iterations'])"
