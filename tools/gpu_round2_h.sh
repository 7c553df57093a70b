#!/bin/bash
mkdir -p gpurun_out
python -m pytest tests/test_gpu_round2.py -m gpu -q --timeout 900 -k "render_thread or adapter" > gpurun_out/pytest_h.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_h.log
grep -E "passed|failed|FAILED|rc=|Error|skipped" gpurun_out/pytest_h.log | tail -8
python bench.py --workload c5 --gpus 1 --steps 2 --warmup 2 > gpurun_out/bench_c5_n1.json 2> gpurun_out/bench_c5_n1.err; echo "c5 n1 rc=$?"; tail -3 gpurun_out/bench_c5_n1.err
python -c "
import json
d=json.loads([l for l in open('gpurun_out/bench_c5_n1.json') if l.startswith('{')][0])
print('c5 n1 value', d['value'], 'ms', d['ms_per_step'], 'e2e', d['e2e']['value'], d['e2e']['ms_per_frame'], d['kernel_ms_per_frame'], d['scene'], d['rays_per_frame'], d['pass_retries'], d.get('cpu_baseline'))
"
