#!/bin/bash
# usage (GPU box, via gpurun): tools/gpu_profile.sh <tag>
# Everything the roofline numbers come from, in one call: a plain run (must exit 0), the ncu launch list of the wavefront
# kernels of one 16-sample pass (duration, DRAM bytes, warp instructions, L2 -> L1 bytes), one --set full capture of the
# level-1 launch of extend / shadow / shade (incoherent rays), and the node-visit / triangle-test counts of the counting build.
T=${1:-r2}
mkdir -p gpurun_out
python tools/profile_frame.py --spp 16 > gpurun_out/${T}_plain.log 2>&1 || { tail -3 gpurun_out/${T}_plain.log; exit 1; }
tail -1 gpurun_out/${T}_plain.log
ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,smsp__inst_executed.sum,smsp__thread_inst_executed.sum,l1tex__m_xbar2l1tex_read_bytes.sum \
    --clock-control none -k regex:k_wf -c 60 --csv --log-file gpurun_out/${T}_launches.csv python tools/profile_frame.py --spp 16 > gpurun_out/${T}_launches.log 2>&1
for k in extend shadow shade; do
  ncu --set full --clock-control none --import-source on -k regex:k_wf_$k -s 1 -c 1 -o gpurun_out/${T}_$k -f python tools/profile_frame.py --spp 16 > gpurun_out/${T}_${k}_ncu.log 2>&1
  tail -1 gpurun_out/${T}_${k}_ncu.log
done
MRT_B200_LIB=$PWD/mobileraytracer_b200/libmobilert_b200_stats.so python tools/traversal_stats.py 2>&1 | tail -2 | tee gpurun_out/${T}_traversal_stats.txt
