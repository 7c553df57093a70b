#!/bin/bash
# usage (here): tools/ncu_read.sh <tag>  -> raw summary + SASS chunk profile of gpurun_out/<tag>.ncu-rep
T=gpurun_out/$1
ncu -i $T.ncu-rep --page raw --csv > ${T}_raw.csv
ncu -i $T.ncu-rep --page source --csv --print-source sass > ${T}_sass.csv 2>/dev/null
python tools/ncu_summary.py ${T}_raw.csv 2>/dev/null | head -${2:-36}
python tools/sass_profile.py ${T}_sass.csv 12
