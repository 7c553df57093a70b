#!/bin/bash
run() { python tools/profile_frame.py --spp 64 | grep -o "'frame_ms': [0-9.]*\|'extend_ms': [0-9.]*\|'shadow_ms': [0-9.]*" | tr '\n' ' '; echo; }
echo -n "base: "; run
for r in 24 28 30; do echo -n "refill=$r: "; MRT_B200_REFILL=$r run; done
for w in 16 18 22; do echo -n "wait=$w any=24: "; MRT_B200_WAIT=$w MRT_B200_WAIT_ANY=24 run; done
for w in 20 22 26 28; do echo -n "wait=20 any=$w: "; MRT_B200_WAIT=20 MRT_B200_WAIT_ANY=$w run; done
for l in 3 5; do echo -n "leaf=$l: "; MRT_B200_LEAF=$l run; done
echo -n "base: "; run
