#!/bin/bash
# 8-GPU call: strong scaling of BASELINE config 4 at N=8 and N=4, and BASELINE config 5 on 8 GPUs (4 tile groups x 2 sample groups)
mkdir -p gpurun_out
run() { # name gpus workload
  python -m torch.distributed.run --nnodes=1 --nproc-per-node $2 --master-addr 127.0.0.1 --master-port 29$((500 + RANDOM % 400)) bench.py --gpus $2 --workload $3 --steps 3 --warmup 3 > gpurun_out/$1.json 2> gpurun_out/$1.err; echo "$1 rc=$?"; grep -E "Error|error" gpurun_out/$1.err | tail -3
}
run bench_c4_n8 8 c4
run bench_c5_n8 8 c5
run bench_c4_n4 4 c4
python -c "
import json
for f in ('bench_c4_n8','bench_c4_n4','bench_c5_n8'):
    try:
        d=json.loads([l for l in open('gpurun_out/%s.json'%f) if l.startswith('{')][0])
        print(f, d['n_gpus'], 'value', round(d['value'],1), 'ms', round(d['ms_per_step'],2), 'e2e', round(d['e2e']['value'],1), round(d['e2e']['ms_per_frame'],2), d['phases_ms_per_frame_rank0'], d['e2e'].get('phases_ms_per_frame_rank0'), 'p+s', round(d['grays_per_s_primary_plus_shadow'],3), d['scene']['build_ms'])
    except Exception as e: print(f, 'ERR', e)
"
