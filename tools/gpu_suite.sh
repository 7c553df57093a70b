#!/bin/bash
# usage (GPU box, via gpurun [--gpus N]): tools/gpu_suite.sh <tag> [N ...]
# The round's acceptance run: the whole GPU test suite, smoke(), the bench line of this library and of the reference arm on
# one GPU, and - for every N given (a box with that many GPUs) - the strong-scaling bench of config 4 and config 5 at N.
T=${1:-r2}; shift
mkdir -p gpurun_out
python -m pytest tests -m gpu -q --timeout 900 > gpurun_out/${T}_pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/${T}_pytest_gpu.log
grep -E "passed|failed|FAILED|rc=" gpurun_out/${T}_pytest_gpu.log | tail -8
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/${T}_smoke.log 2>&1; echo "smoke rc=$?"; tail -3 gpurun_out/${T}_smoke.log
python bench.py --steps 3 --warmup 3 > gpurun_out/${T}_bench_c4_n1.json 2> gpurun_out/${T}_bench_c4_n1.err; echo "bench rc=$?"
python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/${T}_bench_reference_arm.json 2> gpurun_out/${T}_bench_reference_arm.err; echo "reference arm rc=$?"
for N in "$@"; do for W in c4 c5; do
  python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29$((500 + RANDOM % 400)) bench.py --gpus $N --workload $W --steps 3 --warmup 3 \
      > gpurun_out/${T}_bench_${W}_n$N.json 2> gpurun_out/${T}_bench_${W}_n$N.err; echo "$W N=$N rc=$?"
done; done
python - <<PY
import glob, json
for f in sorted(glob.glob("gpurun_out/${T}_bench_c*.json")):
    try:
        d = json.loads([l for l in open(f) if l.startswith("{")][0])
        print(f, "N", d["n_gpus"], "Mrays/s %.0f" % d["value"], "ms %.2f" % d["ms_per_step"], "e2e ms %.2f" % d["e2e"]["ms_per_frame"], d["e2e"].get("phases_ms_per_frame_rank0"))
    except Exception as e:
        print(f, "ERR", e)
PY
