#!/bin/bash
# usage (GPU box with >= max(N) GPUs, via gpurun --gpus N): tools/gpu_scale.sh <tag> <workload:N> [<workload:N> ...]   e.g. r2 c4:8 c5:8 c4:4 c4:2
# Strong-scaling bench lines only (the acceptance run on one GPU is tools/gpu_suite.sh).
T=$1; shift
mkdir -p gpurun_out
for c in "$@"; do
  W=${c%%:*}; N=${c#*:}
  python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29$((500 + RANDOM % 400)) bench.py --gpus $N --workload $W --steps 3 --warmup 3 \
      > gpurun_out/${T}_bench_${W}_n$N.json 2> gpurun_out/${T}_bench_${W}_n$N.err; echo "$W N=$N rc=$?"
done
python - <<PY
import glob, json
for f in sorted(glob.glob("gpurun_out/${T}_bench_c*_n*.json")):
    try:
        d = json.loads([l for l in open(f) if l.startswith("{")][0])
        print(f, "N", d["n_gpus"], "Mrays/s %.0f" % d["value"], "ms %.2f" % d["ms_per_step"], "e2e ms %.2f" % d["e2e"]["ms_per_frame"], d["e2e"].get("phases_ms_per_frame_rank0"), "build %.0f" % d["scene"]["build_ms"])
    except Exception as e:
        print(f, "ERR", e)
PY
