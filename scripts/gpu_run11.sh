mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q -k "two_level or golden" 2>&1 | tail -5
timeout 600 python profiles/precond_sweep.py 2layer_1M full 768 1024 1300 1600 > gpurun_out/precond_1M.log 2>&1; grep -v jacobi gpurun_out/precond_1M.log | cut -c1-330
