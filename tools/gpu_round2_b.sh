#!/bin/bash
# GPU call B: new parity tests, A/B (new / table gather / round-1 library), launch list with DRAM bytes + instruction counts, traversal stats
mkdir -p gpurun_out
python -m pytest tests/test_gpu_round2.py -m gpu -q --timeout 900 > gpurun_out/pytest_b.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_b.log
tail -40 gpurun_out/pytest_b.log
for round in 1 2; do
  echo -n "new   : "; python tools/profile_frame.py --spp 16 --frames 2 | tail -1
  echo -n "gather: "; MRT_B200_TABLE_GATHER=1 python tools/profile_frame.py --spp 16 --frames 2 | tail -1
  echo -n "r1    : "; MRT_B200_LIB=$PWD/mobileraytracer_b200/libmobilert_b200_r1.so python tools/profile_frame.py --spp 16 --frames 2 | tail -1
done 2>&1 | tee gpurun_out/ab_b.txt
ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,smsp__inst_executed.sum,smsp__thread_inst_executed.sum,l1tex__m_xbar2l1tex_read_bytes.sum --clock-control none -c 120 --csv \
    --log-file gpurun_out/launches_b.csv python tools/profile_frame.py --spp 16 > gpurun_out/ncu_b.log 2>&1; tail -1 gpurun_out/ncu_b.log
MRT_B200_LIB=$PWD/mobileraytracer_b200/libmobilert_b200_stats.so python tools/traversal_stats.py 2>&1 | tail -2 | tee gpurun_out/traversal_stats_b.txt
python -m pytest tests/test_gpu_parity.py tests/test_gpu_conference.py -m gpu -q -x --timeout 900 2>&1 | tail -3
