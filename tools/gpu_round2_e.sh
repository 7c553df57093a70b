#!/bin/bash
mkdir -p gpurun_out
python -m pytest tests/test_gpu_round2.py -m gpu -q --timeout 900 -k "conference_pathtracer" > gpurun_out/pytest_e.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_e.log
grep -E "passed|failed|FAILED|rc=|AssertionError" gpurun_out/pytest_e.log | tail -6
for round in 1 2 3; do
  echo -n "new   : "; python tools/profile_frame.py --spp 16 --frames 2 | tail -1
  echo -n "sp    : "; MRT_B200_LIB=$PWD/mobileraytracer_b200/libmobilert_b200_sp.so python tools/profile_frame.py --spp 16 --frames 2 | tail -1
done 2>&1 | tee gpurun_out/ab_e.txt
for v in "20 24" "16 24" "24 28" "20 20" "12 16"; do set -- $v; echo -n "wait $1 any $2: "; MRT_B200_WAIT=$1 MRT_B200_WAIT_ANY=$2 python tools/profile_frame.py --spp 16 --frames 2 | tail -1; done 2>&1 | tee -a gpurun_out/ab_e.txt
for v in 22 24 26 28 30; do echo -n "refill $v: "; MRT_B200_REFILL=$v python tools/profile_frame.py --spp 16 --frames 2 | tail -1; done 2>&1 | tee -a gpurun_out/ab_e.txt
