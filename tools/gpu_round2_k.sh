#!/bin/bash
mkdir -p gpurun_out
python bench.py --steps 3 --warmup 3 > gpurun_out/r2_bench_final.json 2> gpurun_out/r2_bench_final.err; echo "bench rc=$?"; tail -2 gpurun_out/r2_bench_final.err
python -c "
import json
d=json.loads([l for l in open('gpurun_out/r2_bench_final.json') if l.startswith('{')][0])
print(d['value'], d['ms_per_step'], d['e2e']['value'], json.dumps(d['roofline'])[:900]); print(d['roofline_shadow']['bound'], d['roofline_shadow'].get('issue')); print(d['roofline_shade']); print(d['cpu_baseline'])
"
