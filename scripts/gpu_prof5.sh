# round-1 final evidence: launch list of the bench command + full capture of the dominant kernel
mkdir -p gpurun_out
python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/bench_plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -s 4400 -c 1500 --csv --log-file gpurun_out/launches_bench.csv \
    python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_list.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:sell_tma_kernel -s 1500 -c 2 -f -o gpurun_out/sell_tma_full \
    python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_full.log 2>&1
tail -n 1 gpurun_out/bench_plain.log | cut -c1-300
ls -la gpurun_out | head -30
