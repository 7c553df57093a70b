#!/bin/bash
mkdir -p gpurun_out
python -m pytest tests/test_gpu_round2.py -m gpu -q --timeout 900 -k "sort_by_material" > gpurun_out/pytest_f.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_f.log
grep -E "passed|failed|FAILED|rc=|Error" gpurun_out/pytest_f.log | tail -6
for round in 1 2; do
  echo -n "new        : "; python tools/profile_frame.py --spp 16 --frames 2 | tail -1
  echo -n "octant bins: "; MRT_B200_LIB=$PWD/mobileraytracer_b200/libmobilert_b200_oct.so python tools/profile_frame.py --spp 16 --frames 2 | tail -1
  echo -n "sort mat   : "; python tools/profile_frame.py --spp 16 --frames 2 --flags 2 | tail -1
  echo -n "sort prim  : "; MRT_B200_SORT_KEY=primitive python tools/profile_frame.py --spp 16 --frames 2 --flags 2 | tail -1
done 2>&1 | tee gpurun_out/ab_f.txt
