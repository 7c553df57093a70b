#!/bin/bash
# Same-call A/B on the bench frame (GPU box, via gpurun). One script for every variant / tuning experiment:
#   tools/ab.sh [-s SPP] [-r ROUNDS] <case> [<case> ...]
# a case is   base                      the shipped library
#             <name>                    mobileraytracer_b200/libmobilert_b200_<name>.so (tools/build_variant.sh <name> -DFLAG ...)
#             [<name>:]VAR=VALUE[,VAR=VALUE...]   the library (or variant) with environment overrides, e.g.
#                                       MRT_B200_WAIT=16,MRT_B200_WAIT_ANY=24   MRT_B200_REFILL=28   MRT_B200_TABLE_GATHER=1
#                                       flags=2 (profile_frame --flags: MRT_FLAG_SORT_BY_MATERIAL)   MRT_B200_LEAF=3,MRT_B200_FATLEAF=1
# The cases alternate within every round, so clock / box differences hit all of them alike.
SPP=16; ROUNDS=2
while getopts "s:r:" o; do case $o in s) SPP=$OPTARG;; r) ROUNDS=$OPTARG;; esac; done; shift $((OPTIND - 1))
for round in $(seq $ROUNDS); do for c in "$@"; do
  name=${c%%:*}; envs=""; [ "$name" != "$c" ] && envs=${c#*:}
  case "$name" in *=*) envs=$name; name=base;; esac
  lib=""; [ "$name" != base ] && lib=$PWD/mobileraytracer_b200/libmobilert_b200_$name.so
  flags=0; assigns=()
  IFS=',' read -ra kv <<< "$envs"
  for a in "${kv[@]}"; do case "$a" in flags=*) flags=${a#flags=};; ?*) assigns+=("$a");; esac; done
  printf "%-40s " "$c"
  env MRT_B200_LIB=$lib "${assigns[@]}" python tools/profile_frame.py --spp $SPP --frames 2 --flags $flags | tail -1
done; done
