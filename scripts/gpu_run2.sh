mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_gpu.log
tail -n 15 gpurun_out/pytest_gpu.log
python profiles/spmv_sweep.py 2layer_1M > gpurun_out/sweep.log 2>&1; cat gpurun_out/sweep.log | tail -n 3
python profiles/spmv_sweep.py 8layer_10M > gpurun_out/sweep10M.log 2>&1; cat gpurun_out/sweep10M.log | tail -n 3
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 scripts/dist_check.py > gpurun_out/dist2.log 2>&1; echo "dist exit $?" >> gpurun_out/dist2.log
tail -n 20 gpurun_out/dist2.log
