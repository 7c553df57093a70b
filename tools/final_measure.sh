#!/bin/bash
# end-of-round measurement at the state named by $1 (default s42): gpu tests, smoke, bench, reference arm, ncu launch list, ncu --set full of the three frame kernels
S=${1:-s42}
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/${S}_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/${S}_pytest.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > gpurun_out/${S}_smoke.log 2>&1; echo "smoke rc=$?"; tail -4 gpurun_out/${S}_smoke.log
timeout 300 python bench.py --steps 3 --warmup 3 > gpurun_out/${S}_bench.json 2> gpurun_out/${S}_bench.err; echo "bench rc=$?"
timeout 300 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/${S}_ref.json 2> gpurun_out/${S}_ref.err; echo "ref rc=$?"
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/${S}_launches.csv python tools/profile_frame.py --spp 16 > gpurun_out/${S}_launch_ncu.log 2>&1; echo "ncu rc=$?"
timeout 400 tools/ncu_kernel.sh k_wf_extend ${S}_extend
timeout 300 tools/ncu_kernel.sh k_wf_shadow ${S}_shadow
cat gpurun_out/${S}_bench.json
