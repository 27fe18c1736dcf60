mkdir -p gpurun_out
N=${1:-2}
PADNE_B200_KTIME=1 timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus $N --steps 1 --warmup 3 --no-cpu-baseline > gpurun_out/ktime$N.log 2>&1; echo "bench exit $?" >> gpurun_out/ktime$N.log
grep -n "ktime rank" gpurun_out/ktime$N.log | tail -4
tail -40 gpurun_out/ktime$N.log | grep -E "ktime|n=|exit"
