#!/bin/bash
# usage (GPU box, via gpurun): tools/gpu_suite.sh <tag>
# The round's acceptance run on one GPU: the whole GPU test suite, smoke(), the bench line of this library and of the reference arm.
# (strong scaling on several GPUs: tools/gpu_scale.sh)
T=${1:-r2}
mkdir -p gpurun_out
python -m pytest tests -m gpu -q --timeout 900 > gpurun_out/${T}_pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/${T}_pytest_gpu.log
grep -E "passed|failed|FAILED|rc=" gpurun_out/${T}_pytest_gpu.log | tail -8
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/${T}_smoke.log 2>&1; echo "smoke rc=$?"; tail -3 gpurun_out/${T}_smoke.log
python bench.py --steps 3 --warmup 3 > gpurun_out/${T}_bench_c4_n1.json 2> gpurun_out/${T}_bench_c4_n1.err; echo "bench rc=$?"
python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/${T}_bench_reference_arm.json 2> gpurun_out/${T}_bench_reference_arm.err; echo "reference arm rc=$?"
python - <<PY
import glob, json
for f in sorted(glob.glob("gpurun_out/${T}_bench_c*.json")):
    try:
        d = json.loads([l for l in open(f) if l.startswith("{")][0])
        print(f, "N", d["n_gpus"], "Mrays/s %.0f" % d["value"], "ms %.2f" % d["ms_per_step"], "e2e ms %.2f" % d["e2e"]["ms_per_frame"], d["e2e"].get("phases_ms_per_frame_rank0"))
    except Exception as e:
        print(f, "ERR", e)
PY
