#!/bin/bash
# usage (GPU box): tools/ncu_traffic.sh <tag>  -> DRAM bytes + duration of every k_wf_extend / k_wf_shadow launch of one 16-sample pass
python tools/profile_frame.py --spp 16 > gpurun_out/$1_plain.log 2>&1 && \
ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum --clock-control none -k regex:"k_wf_extend|k_wf_shadow" -c 12 --csv \
    --log-file gpurun_out/$1.csv python tools/profile_frame.py --spp 16 > gpurun_out/$1_ncu.log 2>&1
tail -1 gpurun_out/$1_plain.log
