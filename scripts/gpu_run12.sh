mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -8
timeout 600 python profiles/precond_sweep.py 2layer_1M full 968 1250 1458 1700 > gpurun_out/precond_1M.log 2>&1; grep -v jacobi gpurun_out/precond_1M.log | cut -c1-330
timeout 600 python profiles/precond_sweep.py 4layer_regulator_4M full 512 1024 1600 > gpurun_out/precond_4M.log 2>&1; cut -c1-420 gpurun_out/precond_4M.log
