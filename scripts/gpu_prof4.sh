mkdir -p gpurun_out
python profiles/precond_sweep.py 2layer_1M full 1024 > gpurun_out/plain.log 2>&1 &&
ncu --set full --clock-control none --cache-control none --import-source on -k regex:"cg_update_kernel" -s 8600 -c 1 -f -o gpurun_out/cgupd_tl \
    python profiles/precond_sweep.py 2layer_1M full 1024 > gpurun_out/ncu_a.log 2>&1
ncu --set full --clock-control none --cache-control none --import-source on -k regex:"coarse_apply" -s 100 -c 1 -f -o gpurun_out/capply \
    python profiles/precond_sweep.py 2layer_1M full 1024 > gpurun_out/ncu_b.log 2>&1
tail -n 2 gpurun_out/ncu_a.log gpurun_out/ncu_b.log
