mkdir -p gpurun_out
timeout 300 python profiles/one_solve.py 2>&1 | tail -2
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 320 --csv --log-file gpurun_out/setup_launches.csv python profiles/one_solve.py > gpurun_out/ncu_setup.log 2>&1
tail -2 gpurun_out/ncu_setup.log
