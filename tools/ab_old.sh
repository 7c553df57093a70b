#!/bin/bash
# same-call A/B: shipped library vs the "old" variant (tools/build_variant.sh old at the previous commit)
run() { python tools/profile_frame.py --spp 64 | grep -o "'frame_ms': [0-9.]*\|'extend_ms': [0-9.]*\|'shadow_ms': [0-9.]*" | tr '\n' ' '; echo; }
for round in 1 2 3; do
  echo -n "old: "; MRT_B200_LIB=$PWD/mobileraytracer_b200/libmobilert_b200_old.so run
  echo -n "new: "; run
done
python -m pytest tests/test_gpu_conference.py -m gpu -x -q -k "incoherent or primary or pathtracer or full_size or determin" 2>&1 | tail -3
