#!/bin/bash
for l in 2 3 4 5 7; do echo -n "leaf=$l "; MRT_B200_LEAF=$l python tools/profile_frame.py --spp 4 | grep -o "'extend_ms': [0-9.]*, 'shadow_ms': [0-9.]*\|'bvh_nodes': [0-9]*" | tr '\n' ' '; MRT_B200_LEAF=$l MRT_B200_LIB=$PWD/mobileraytracer_b200/libmobilert_b200_stats.so python tools/traversal_stats.py | grep -o "nodes/ray [0-9.]*  tris/ray [0-9.]*"; done
