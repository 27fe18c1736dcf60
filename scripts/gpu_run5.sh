mkdir -p gpurun_out
timeout 300 python profiles/precond_sweep.py 2layer_1M 200 64 256 > gpurun_out/precond_small.log 2>&1; echo "small exit $?" >> gpurun_out/precond_small.log
cat gpurun_out/precond_small.log | tail -n 8
timeout 600 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_gpu.log
tail -n 12 gpurun_out/pytest_gpu.log
timeout 600 python profiles/precond_sweep.py 2layer_1M full 256 512 1024 1536 > gpurun_out/precond_1M.log 2>&1; echo "1M exit $?" >> gpurun_out/precond_1M.log
cat gpurun_out/precond_1M.log | tail -n 8
