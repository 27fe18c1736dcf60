# 8-GPU record: three sizes of the refinement sweep (stack_16layer, n per layer; 2.4M / 24M / 50M DOF)
N=${1:-8}
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517"
out=gpurun_out/sweep_${N}gpu.jsonl; : > $out
for n in 387 1225 1768; do
  timeout 600 $TR bench.py --gpus $N --grid $n --steps 3 --warmup 3 --no-named-configs --no-cpu-baseline 2>gpurun_out/sweep_err.log | grep '^{' >> $out; echo "n=$n exit ${PIPESTATUS[0]}"
done
python - <<PY
import json
for line in open("$out"):
    d = json.loads(line); c = d["config"]
    print(c["dof"], "DOF", f'{d["value"] / 1e6:.1f} MDOF/s', f'{d["ms_per_step"]:.1f} ms', c["iterations"], "iters", round(c["us_per_iteration"], 1), "us/it", "res", c["residual_norm"])
PY
