mkdir -p gpurun_out
timeout 600 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_gpu.log
tail -n 4 gpurun_out/pytest_gpu.log
timeout 600 python profiles/precond_sweep.py 2layer_1M full 512 1024 > gpurun_out/precond_1M.log 2>&1; tail -n 3 gpurun_out/precond_1M.log
timeout 600 python profiles/precond_sweep.py 8layer_10M full 1024 1536 > gpurun_out/precond_10M.log 2>&1; tail -n 3 gpurun_out/precond_10M.log
timeout 600 python bench.py --steps 5 --warmup 3 > gpurun_out/bench.log 2>&1; echo "bench exit $?" >> gpurun_out/bench.log; tail -n 2 gpurun_out/bench.log
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 scripts/dist_check.py > gpurun_out/dist2.log 2>&1; echo "dist exit $?" >> gpurun_out/dist2.log
grep -E '^\{|exit' gpurun_out/dist2.log
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus 2 --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/bench2.log 2>&1; echo "bench2 exit $?" >> gpurun_out/bench2.log
tail -n 2 gpurun_out/bench2.log
