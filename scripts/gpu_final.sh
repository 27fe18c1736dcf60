# final verification of the committed tree on one GPU (+ ncu launch list / SpMV capture when PROF=1)
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -3
timeout 300 python __graft_entry__.py --smoke 2>&1 | tail -2
timeout 900 python bench.py --steps 5 --warmup 3 > gpurun_out/bench_final.log 2>&1; echo "bench exit $?"; grep '^{' gpurun_out/bench_final.log | cut -c1-1500
timeout 900 python bench.py --impl reference --steps 1 --warmup 0 > gpurun_out/bench_ref.log 2>&1; echo "ref exit $?"; grep '^{' gpurun_out/bench_ref.log | cut -c1-600
if [ "$PROF" = "1" ]; then
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 700 --csv --log-file gpurun_out/launches_final.csv python bench.py --steps 1 --warmup 1 --no-cpu-baseline > gpurun_out/ncu_final_list.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:sell_tma -s 30 -c 1 -o gpurun_out/sell_final -f python bench.py --steps 1 --warmup 1 --no-cpu-baseline > gpurun_out/ncu_final_full.log 2>&1
tail -1 gpurun_out/ncu_final_full.log
fi
