#!/bin/bash
# usage (GPU box): tools/ab.sh <spp> <variant> [<variant> ...]   ("base" = the shipped library); two alternating rounds
SPP=$1; shift
for round in 1 2; do for v in "$@"; do
  if [ "$v" = base ]; then L=""; else L=$PWD/mobileraytracer_b200/libmobilert_b200_$v.so; fi
  echo -n "$v: "; MRT_B200_LIB=$L python tools/profile_frame.py --spp $SPP | grep -o "'frame_ms': [0-9.]*\|'extend_ms': [0-9.]*\|'shadow_ms': [0-9.]*\|'shade_ms': [0-9.]*\|'other_ms': [0-9.]*" | tr '\n' ' '; echo
done; done
