# round 2: the 1-GPU record -- tests, smoke, both bench arms, ncu launch list + full captures of the dominant kernels
mkdir -p gpurun_out
timeout 300 python __graft_entry__.py --smoke 2>&1 | tail -1
( time timeout 900 python bench.py --steps 5 --warmup 3 > gpurun_out/r2_bench_1gpu.log 2>&1 ) 2>&1 | grep real; echo "bench exit $?"; grep '^{' gpurun_out/r2_bench_1gpu.log | cut -c1-3000
( time timeout 900 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/r2_bench_ref.log 2>&1 ) 2>&1 | grep real; grep '^{' gpurun_out/r2_bench_ref.log | cut -c1-1200
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 2500 --csv --log-file gpurun_out/r2_launches_2layer_1M.csv python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-named-configs > gpurun_out/ncu_list.log 2>&1; tail -1 gpurun_out/ncu_list.log | cut -c1-200
ONE_SOLVE_ITERS=64 timeout 600 ncu --set full --clock-control none --import-source on -k regex:sell_tma_kernel -s 40 -c 1 -o gpurun_out/r2_spmv_cg -f python profiles/one_solve.py > gpurun_out/ncu_a.log 2>&1; tail -1 gpurun_out/ncu_a.log
ONE_SOLVE_ITERS=64 timeout 600 ncu --set full --clock-control none --import-source on -k regex:cg_update_tl_kernel -s 40 -c 1 -o gpurun_out/r2_update_tl -f python profiles/one_solve.py > gpurun_out/ncu_b.log 2>&1; tail -1 gpurun_out/ncu_b.log
