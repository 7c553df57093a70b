for round in 1 2; do for ov in 0 1; do
  echo -n "overlap=$ov: "; MRT_B200_OVERLAP=$ov python tools/profile_frame.py --spp 64 | grep -o "'frame_ms': [0-9.]*\|'extend_ms': [0-9.]*\|'shadow_ms': [0-9.]*\|'shade_ms': [0-9.]*\|'other_ms': [0-9.]*" | tr '\n' ' '; echo
done; done
python -m pytest tests/test_gpu_conference.py -m gpu -x -q 2>&1 | tail -3
