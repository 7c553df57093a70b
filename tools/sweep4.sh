#!/bin/bash
run() { python tools/profile_frame.py --spp 64 | grep -o "'frame_ms': [0-9.]*\|'extend_ms': [0-9.]*\|'shadow_ms': [0-9.]*" | tr '\n' ' '; echo; }
echo -n "base: "; run
echo -n "no node prefetch: "; MRT_B200_LIB=$PWD/mobileraytracer_b200/libmobilert_b200_nopf.so run
for r in 16 22 26 29; do for w in 4 8 12 20; do echo -n "refill=$r wait=$w: "; MRT_B200_REFILL=$r MRT_B200_WAIT=$w run; done; done
echo -n "base: "; run
