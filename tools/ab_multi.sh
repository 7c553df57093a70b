#!/bin/bash
# usage (GPU box): tools/ab_multi.sh <variant> ...   ("base" = shipped library); three alternating rounds at 64 spp
run() { python tools/profile_frame.py --spp 64 | grep -o "'frame_ms': [0-9.]*\|'extend_ms': [0-9.]*\|'shadow_ms': [0-9.]*\|'shade_ms': [0-9.]*" | tr '\n' ' '; echo; }
for round in 1 2 3; do for v in "$@"; do
  if [ "$v" = base ]; then L=""; else L=$PWD/mobileraytracer_b200/libmobilert_b200_$v.so; fi
  echo -n "$v: "; MRT_B200_LIB=$L run
done; done
