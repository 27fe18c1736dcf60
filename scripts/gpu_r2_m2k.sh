N=${1:-2}
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29533"
PADNE_B200_KTIME=1 timeout 600 $TR bench.py --gpus $N --steps 1 --warmup 3 --no-cpu-baseline --no-named-configs 2>&1 | grep -A6 "ktime rank 0" | head -24
