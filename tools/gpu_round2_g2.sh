#!/bin/bash
# 2-GPU call: the NCCL frame against the single-rank frame, and the strong-scaling bench at N=2 and N=1 on the same box
mkdir -p gpurun_out
python -m pytest tests/test_gpu_round2.py -m gpu -q --timeout 900 -k "two_rank" > gpurun_out/pytest_g2.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_g2.log
grep -E "passed|failed|FAILED|rc=|Error|skipped" gpurun_out/pytest_g2.log | tail -6
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29555 bench.py --gpus 2 --steps 3 --warmup 3 > gpurun_out/bench_n2.json 2> gpurun_out/bench_n2.err; echo "n2 rc=$?"; tail -2 gpurun_out/bench_n2.err
python bench.py --gpus 1 --steps 3 --warmup 3 > gpurun_out/bench_n1.json 2> gpurun_out/bench_n1.err; echo "n1 rc=$?"; tail -2 gpurun_out/bench_n1.err
python -c "
import json
for f in ('gpurun_out/bench_n1.json','gpurun_out/bench_n2.json'):
    d=json.load(open(f)); print(f, d['n_gpus'], 'value', d['value'], 'ms', d['ms_per_step'], 'e2e', d['e2e']['value'], d['e2e']['ms_per_frame'], d['phases_ms_per_frame_rank0'], d['e2e'].get('phases_ms_per_frame_rank0'), d['grays_per_s_primary_plus_shadow'])
"
