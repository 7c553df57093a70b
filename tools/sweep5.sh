#!/bin/bash
export MRT_B200_LIB=$PWD/mobileraytracer_b200/libmobilert_b200_nopf.so
run() { python tools/profile_frame.py --spp 64 | grep -o "'frame_ms': [0-9.]*\|'extend_ms': [0-9.]*\|'shadow_ms': [0-9.]*" | tr '\n' ' '; echo; }
for r in 22 26 29; do for w in 16 20 24 28 33; do echo -n "nopf refill=$r wait=$w: "; MRT_B200_REFILL=$r MRT_B200_WAIT=$w run; done; done
