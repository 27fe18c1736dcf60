mkdir -p gpurun_out
python profiles/precond_sweep.py 2layer_1M full 1024 > gpurun_out/plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none --cache-control none -k regex:"cg_update|coarse|sell_tma|cg_init" -s 17400 -c 60 --csv --log-file gpurun_out/launches_tl.csv \
    python profiles/precond_sweep.py 2layer_1M full 1024 > gpurun_out/ncu_tl.log 2>&1
tail -n 3 gpurun_out/plain.log
grep -c sell_tma gpurun_out/launches_tl.csv
