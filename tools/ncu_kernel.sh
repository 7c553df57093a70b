#!/bin/bash
# usage (on the GPU box, via gpurun): tools/ncu_kernel.sh <kernel regex> <tag> [profile_frame args]
# plain run first (must exit 0), then one ncu --set full capture of 2 launches -> gpurun_out/<tag>.ncu-rep
K=$1; TAG=$2; shift 2
python tools/profile_frame.py --spp 2 "$@" > gpurun_out/${TAG}_plain.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:$K -s 1 -c 2 -o gpurun_out/$TAG -f python tools/profile_frame.py --spp 2 "$@" > gpurun_out/${TAG}_ncu.log 2>&1
tail -1 gpurun_out/${TAG}_plain.log; tail -1 gpurun_out/${TAG}_ncu.log
