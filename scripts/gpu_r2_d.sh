mkdir -p gpurun_out
ONE_SOLVE_ITERS=64 timeout 600 ncu --set full --clock-control none --import-source on -k regex:sell_tma_kernel -s 40 -c 1 -o gpurun_out/r2_ring_cg2 -f python profiles/one_solve.py > gpurun_out/ncu_ring_cg2.log 2>&1; tail -1 gpurun_out/ncu_ring_cg2.log
ONE_SOLVE_ITERS=64 timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_v2.csv python profiles/one_solve.py > gpurun_out/ncu_list.log 2>&1; tail -1 gpurun_out/ncu_list.log
