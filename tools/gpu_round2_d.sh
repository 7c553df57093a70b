#!/bin/bash
mkdir -p gpurun_out
python -m pytest tests -m gpu -q --timeout 900 > gpurun_out/pytest_d.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_d.log
grep -E "passed|failed|FAILED|rc=" gpurun_out/pytest_d.log | tail -12
for round in 1 2; do
  echo -n "new   : "; python tools/profile_frame.py --spp 16 --frames 2 | tail -1
  echo -n "gather: "; MRT_B200_TABLE_GATHER=1 python tools/profile_frame.py --spp 16 --frames 2 | tail -1
  echo -n "sb3   : "; MRT_B200_LIB=$PWD/mobileraytracer_b200/libmobilert_b200_sb3.so python tools/profile_frame.py --spp 16 --frames 2 | tail -1
  echo -n "sb1   : "; MRT_B200_LIB=$PWD/mobileraytracer_b200/libmobilert_b200_sb1.so python tools/profile_frame.py --spp 16 --frames 2 | tail -1
  echo -n "r1    : "; MRT_B200_LIB=$PWD/mobileraytracer_b200/libmobilert_b200_r1.so python tools/profile_frame.py --spp 16 --frames 2 | tail -1
done 2>&1 | tee gpurun_out/ab_d.txt
ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,smsp__inst_executed.sum,smsp__thread_inst_executed.sum,l1tex__m_xbar2l1tex_read_bytes.sum --clock-control none -k regex:k_wf -c 60 --csv \
    --log-file gpurun_out/launches_d.csv python tools/profile_frame.py --spp 16 > gpurun_out/ncu_d.log 2>&1; tail -1 gpurun_out/ncu_d.log
python bench.py --steps 2 --warmup 2 --no-cpu-baseline > gpurun_out/bench_d.json 2> gpurun_out/bench_d.err; echo "bench rc=$?"; tail -3 gpurun_out/bench_d.err
python -c "
import json; d=json.load(open('gpurun_out/bench_d.json'))
for k in ['value','ms_per_step','kernel_ms_per_frame','grays_per_s_primary_plus_shadow','gpu_launches','phases_ms_per_frame_rank0']: print(k, d.get(k))
print('e2e', d['e2e']['value'], d['e2e']['ms_per_frame'])
"
