# launch list of the bench command + one full capture of the dominant kernel (run after a clean bench)
mkdir -p gpurun_out
python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/bench_pre.log 2>&1 || exit 1
ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/launches.csv \
    python bench.py --steps 1 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_list.log 2>&1
python profiles/spmv_sweep.py 2layer_1M > gpurun_out/sweep.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:spmv_kernel -s 200 -c 2 -f -o gpurun_out/spmv_full \
    python profiles/spmv_sweep.py 2layer_1M > gpurun_out/ncu_full.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:cg_update -s 200 -c 1 -f -o gpurun_out/cgupd_full \
    python profiles/spmv_sweep.py 2layer_1M > gpurun_out/ncu_full2.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:row_kernel -c 1 -f -o gpurun_out/row_full \
    python profiles/spmv_raw.py 2layer_1M > gpurun_out/ncu_full3.log 2>&1
python profiles/spmv_sweep.py 8layer_10M > gpurun_out/sweep10M.log 2>&1
cat gpurun_out/sweep.log gpurun_out/sweep10M.log
ls -la gpurun_out
