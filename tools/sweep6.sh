#!/bin/bash
run() { python tools/profile_frame.py --spp 64 | grep -o "'frame_ms': [0-9.]*\|'extend_ms': [0-9.]*\|'shadow_ms': [0-9.]*\|'shade_ms': [0-9.]*\|'other_ms': [0-9.]*" | tr '\n' ' '; echo; }
for round in 1 2; do
echo -n "base: "; run
echo -n "no queue prefetch: "; MRT_B200_LIB=$PWD/mobileraytracer_b200/libmobilert_b200_noqpf.so run
echo -n "leaf=2: "; MRT_B200_LEAF=2 run
echo -n "leaf=4: "; MRT_B200_LEAF=4 run
echo -n "fat leaf=4: "; MRT_B200_FATLEAF=1 MRT_B200_LEAF=4 run
echo -n "fat leaf=6: "; MRT_B200_FATLEAF=1 MRT_B200_LEAF=6 run
done
python -m pytest tests -m gpu -x -q 2>&1 | tail -3
