#!/bin/bash
# same-call A/B: shipped library vs the "old" variant, plain and with fat leaves of several sizes
run() { python tools/profile_frame.py --spp 64 | grep -o "'frame_ms': [0-9.]*\|'extend_ms': [0-9.]*\|'shadow_ms': [0-9.]*\|'bvh_nodes': [0-9]*" | tr '\n' ' '; echo; }
for round in 1 2; do
  echo -n "old: "; MRT_B200_LIB=$PWD/mobileraytracer_b200/libmobilert_b200_old.so run
  echo -n "new: "; run
  for l in 3 4 5 6; do echo -n "new fat leaf=$l: "; MRT_B200_FATLEAF=1 MRT_B200_LEAF=$l run; done
  echo -n "old fat leaf=4: "; MRT_B200_LIB=$PWD/mobileraytracer_b200/libmobilert_b200_old.so MRT_B200_FATLEAF=1 MRT_B200_LEAF=4 run
done
python -m pytest tests/test_gpu_conference.py -m gpu -x -q -k "incoherent or primary or pathtracer" 2>&1 | tail -3
