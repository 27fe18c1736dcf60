mkdir -p gpurun_out
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 scripts/dist_check.py > gpurun_out/dist2.log 2>&1; echo "dist exit $?" >> gpurun_out/dist2.log
grep -E '^\{|exit|rror' gpurun_out/dist2.log | cut -c1-700
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus 2 --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/bench2.log 2>&1; echo "bench2 exit $?" >> gpurun_out/bench2.log
tail -n 2 gpurun_out/bench2.log | cut -c1-900
PADNE_B200_COMM=nccl timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29513 bench.py --gpus 2 --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/bench2_nccl.log 2>&1; echo "bench2n exit $?" >> gpurun_out/bench2_nccl.log
tail -n 2 gpurun_out/bench2_nccl.log | cut -c1-600
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29514 bench.py --gpus 2 --steps 2 --warmup 3 --no-cpu-baseline --config 8layer_10M > gpurun_out/bench2_10M.log 2>&1; echo "bench2_10M exit $?" >> gpurun_out/bench2_10M.log
tail -n 2 gpurun_out/bench2_10M.log | cut -c1-900
