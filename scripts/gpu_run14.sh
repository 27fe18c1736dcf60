mkdir -p gpurun_out
N=${1:-2}
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29513 profiles/dist_probe.py stack_$((2*N))layer 640 > gpurun_out/probe$N.log 2>&1; echo "probe exit $?" >> gpurun_out/probe$N.log
grep -E '^\{|exit' gpurun_out/probe$N.log
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 scripts/dist_check.py 2layer_1M 300 4layer_regulator_4M 120 8layer_10M 300 > gpurun_out/dist$N.log 2>&1; echo "dist exit $?" >> gpurun_out/dist$N.log
grep -E '^\{|exit|rror' gpurun_out/dist$N.log | cut -c1-600
PADNE_B200_TRACE=1 PADNE_B200_PEER_STATS=1 timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus $N --steps 2 --warmup 3 --no-cpu-baseline --config 8layer_10M > gpurun_out/trace$N.log 2>&1; echo "bench exit $?" >> gpurun_out/trace$N.log
grep -E "trace rank 0|exit|exchanges" gpurun_out/trace$N.log | tail -5
