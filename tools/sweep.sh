#!/bin/bash
# tuning sweep of the traversal scheduling thresholds on the bench workload (4 spp)
for p in ${POSTPONE:-1 6 10 14 18 24}; do for r in ${REFILL:-22}; do
  echo -n "postpone=$p refill=$r  "; MRT_B200_POSTPONE=$p MRT_B200_REFILL=$r python tools/profile_frame.py --spp 4 | grep -o "'extend_ms': [0-9.]*, 'shadow_ms': [0-9.]*"
done; done
