mkdir -p gpurun_out
ONE_SOLVE_ITERS=64 timeout 600 ncu --set full --clock-control none --cache-control none --import-source on -k regex:cg_update_tl -s 20 -c 1 -o gpurun_out/cgupd_tl2 -f python profiles/one_solve.py > gpurun_out/ncu_c.log 2>&1
tail -2 gpurun_out/ncu_c.log
