python bench.py --steps 3 --warmup 3 > gpurun_out/s40_bench.json 2> gpurun_out/s40_bench.err; echo "bench rc=$?"
python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/s40_ref.json 2> gpurun_out/s40_ref.err; echo "ref rc=$?"
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/s40_launches.csv python tools/profile_frame.py --spp 16 > gpurun_out/s40_launch_ncu.log 2>&1; echo "ncu rc=$?"
python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')"
tools/ncu_traffic.sh s40_traffic
tools/ncu_kernel.sh k_wf_extend s40_extend
tools/ncu_kernel.sh k_wf_shadow s40_shadow
tools/ncu_kernel.sh k_wf_shade s40_shade
