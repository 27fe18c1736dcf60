mkdir -p gpurun_out
ONE_SOLVE_ITERS=64 timeout 300 python profiles/one_solve.py 2>&1 | tail -1
ONE_SOLVE_ITERS=64 timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none --cache-control none -c 360 --csv --log-file gpurun_out/warm_launches.csv python profiles/one_solve.py > gpurun_out/ncu_warm.log 2>&1
tail -2 gpurun_out/ncu_warm.log
