mkdir -p gpurun_out
timeout 600 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_gpu.log
tail -n 12 gpurun_out/pytest_gpu.log
timeout 600 python profiles/precond_sweep.py 2layer_1M full 1024 > gpurun_out/precond_1M.log 2>&1; tail -n 2 gpurun_out/precond_1M.log
