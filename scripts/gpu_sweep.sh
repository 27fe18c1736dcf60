# BASELINE.json configs[4]: mesh-refinement sweep of the 2-layer plane board, 0.1M -> 50M DOF (DOF ~ 2 n^2).
#   gpurun --timeout 1500 -- 'bash scripts/gpu_sweep.sh'            one GPU
#   gpurun --gpus 8 --timeout 1500 -- 'bash scripts/gpu_sweep.sh 8'   partitioned stack_16layer, n per layer
# One JSON line per size in gpurun_out/sweep_<N>gpu.jsonl; the CPU baseline leg runs once, on the first size.
N=${1:-1}
mkdir -p gpurun_out
out=gpurun_out/sweep_${N}gpu.jsonl
: > $out
first=1
for n in 224 387 708 1225 2236 3873 5000; do
  [ "$N" -gt 1 ] && [ $n -gt 2236 ] && break          # per-GPU share of the stack stays under 10M DOF
  extra="--no-cpu-baseline"; [ $first = 1 ] && extra=""; first=0
  if [ "$N" = 1 ]; then
    timeout 900 python bench.py --grid $n --steps 3 --warmup 3 --no-named-configs $extra 2>gpurun_out/sweep_err.log | grep '^{' >> $out
  else
    timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517 \
      bench.py --gpus $N --grid $n --steps 3 --warmup 3 --no-named-configs $extra 2>gpurun_out/sweep_err.log | grep '^{' >> $out
  fi
  echo "n=$n exit ${PIPESTATUS[0]}"
done
python - <<'PY'
import json, sys, glob
for p in sorted(glob.glob("gpurun_out/sweep_*gpu.jsonl")):
    for line in open(p):
        d = json.loads(line)
        print(p, d["config"].get("dof"), "DOF", f'{d["value"] / 1e6:.1f} MDOF/s', f'{d["ms_per_step"]:.1f} ms',
              "spmv frac", d["roofline"]["frac"])
PY
