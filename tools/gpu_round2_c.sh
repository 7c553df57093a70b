#!/bin/bash
mkdir -p gpurun_out
python -m pytest tests/test_gpu_round2.py -m gpu -q --timeout 900 -k "conference_pathtracer or depthmap_full or adapter" > gpurun_out/pytest_c.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_c.log
tail -12 gpurun_out/pytest_c.log
for round in 1 2; do
  echo -n "new   : "; python tools/profile_frame.py --spp 16 --frames 2 | tail -1
  echo -n "gather: "; MRT_B200_TABLE_GATHER=1 python tools/profile_frame.py --spp 16 --frames 2 | tail -1
  echo -n "sb3   : "; MRT_B200_LIB=$PWD/mobileraytracer_b200/libmobilert_b200_sb3.so python tools/profile_frame.py --spp 16 --frames 2 | tail -1
  echo -n "sb2   : "; MRT_B200_LIB=$PWD/mobileraytracer_b200/libmobilert_b200_sb2.so python tools/profile_frame.py --spp 16 --frames 2 | tail -1
done 2>&1 | tee gpurun_out/ab_c.txt
ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,smsp__inst_executed.sum,smsp__thread_inst_executed.sum,l1tex__m_xbar2l1tex_read_bytes.sum --clock-control none -k regex:k_wf -c 60 --csv \
    --log-file gpurun_out/launches_c.csv python tools/profile_frame.py --spp 16 > gpurun_out/ncu_c.log 2>&1; tail -1 gpurun_out/ncu_c.log
ncu --set full --clock-control none --import-source on -k regex:k_wf_shade -s 1 -c 1 -o gpurun_out/shade_c -f python tools/profile_frame.py --spp 16 > gpurun_out/ncu_shade_c.log 2>&1; tail -1 gpurun_out/ncu_shade_c.log
ncu --set full --clock-control none --import-source on -k regex:k_wf_extend -s 1 -c 1 -o gpurun_out/extend_c -f python tools/profile_frame.py --spp 16 > gpurun_out/ncu_extend_c.log 2>&1; tail -1 gpurun_out/ncu_extend_c.log
