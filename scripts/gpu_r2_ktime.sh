# per-kernel times of the partitioned PCG iteration with and without the rank-local level (event after every launch)
N=${1:-2}
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29533"
for ll in 0 1; do
echo "== LOCAL_LEVEL=$ll"
PADNE_B200_LOCAL_LEVEL=$ll PADNE_B200_KTIME=1 PADNE_B200_PEER_STATS=1 timeout 600 $TR bench.py --gpus $N --steps 3 --warmup 3 --no-cpu-baseline --no-named-configs 2>&1 | grep -E "ktime rank 0|mean|event intervals|^\{" | grep -v "rank [1-9]" | cut -c1-400 | head -24
done
