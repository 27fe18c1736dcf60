mkdir -p gpurun_out
: > gpurun_out/variants2.log
for lib in padne_b200/_lib/libpadne_b200.so padne_b200/_lib/libpadne_b200_w8.so padne_b200/_lib/libpadne_b200_tw16w8.so padne_b200/_lib/libpadne_b200_tw12m3w8.so padne_b200/_lib/libpadne_b200_tw8m4w8.so; do
  for cfg in 2layer_1M 8layer_10M; do
    PADNE_B200_LIB=$PWD/$lib timeout 300 python profiles/spmv_sweep.py $cfg 2>&1 | tail -n 1 | sed "s|^|$(basename $lib) |" >> gpurun_out/variants2.log
  done
done
cat gpurun_out/variants2.log | python -c "
import sys, json
for l in sys.stdin:
    lib, js = l.split(' ', 1)
    try: d = json.loads(js)
    except Exception: print(lib, js[:200]); continue
    print(f\"{lib:28s} {d['config']:11s} it={d['iterations']} us/it={d['us_per_iteration']:.1f} warm={d['warm_us']:.1f} cold={d['cold_us']:.1f} insitu={d['in_situ_us']:.1f}us {d['in_situ_GBs']:.0f}GB/s\")
"
