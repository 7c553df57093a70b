#!/bin/bash
mkdir -p gpurun_out
bash tools/ab.sh -r 2 base 2>&1 | tee gpurun_out/ab_n.txt
python tools/run_configs.py > gpurun_out/r2_configs_1_3.jsonl 2> gpurun_out/r2_configs.err; cat gpurun_out/r2_configs_1_3.jsonl | cut -c1-330
python -m pytest tests -m gpu -q -x --timeout 900 2>&1 | tail -3
