#!/bin/bash
# GPU call: the whole GPU test suite, a short bench line and a same-box A/B of this library against the round-1 library
mkdir -p gpurun_out
python -m pytest tests -m gpu -q -x --timeout 900 > gpurun_out/pytest_a.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_a.log
tail -15 gpurun_out/pytest_a.log
for round in 1 2; do
  echo -n "new: "; python tools/profile_frame.py --spp 16 --frames 2 | tail -1
  echo -n "r1 : "; MRT_B200_LIB=$PWD/mobileraytracer_b200/libmobilert_b200_r1.so python tools/profile_frame.py --spp 16 --frames 2 | tail -1
done 2>&1 | tee gpurun_out/ab_r1_vs_new.txt
python bench.py --steps 2 --warmup 2 --no-cpu-baseline > gpurun_out/bench_a.json 2> gpurun_out/bench_a.err; echo "bench rc=$?"; tail -3 gpurun_out/bench_a.err
python -c "
import json; d=json.load(open('gpurun_out/bench_a.json'))
for k in ['value','ms_per_step','kernel_ms_per_frame','rays_per_frame','grays_per_s_primary_plus_shadow','gpu_launches','phases_ms_per_frame_rank0','gather_peaks','pass_retries']: print(k, d.get(k))
print('e2e', d['e2e']['value'], d['e2e']['ms_per_frame'])
"
