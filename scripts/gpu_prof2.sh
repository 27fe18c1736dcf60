mkdir -p gpurun_out
python profiles/spmv_sweep.py 2layer_1M > gpurun_out/sweep.log 2>&1 || exit 1
ncu --set full --clock-control none --cache-control none --import-source on -k regex:sell_spmv -s 4300 -c 2 -f -o gpurun_out/sell_warm \
    python profiles/spmv_sweep.py 2layer_1M > gpurun_out/ncu_sell.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:sell_spmv -s 4300 -c 1 -f -o gpurun_out/sell_cold \
    python profiles/spmv_sweep.py 2layer_1M > gpurun_out/ncu_sell2.log 2>&1
tail -n 3 gpurun_out/ncu_sell.log
