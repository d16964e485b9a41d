#!/bin/bash
# hand-over of long single rays (CRT_OPT_RAY_DEFER) on top of the packet hand-over: parity, then ranks' shares of 8
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q -k "packet_defer or wide_bounces or tail_hint" > gpurun_out/p3_pytest.log 2>&1
echo "pytest rc $?"; tail -4 gpurun_out/p3_pytest.log
for after in 0 128 192 256; do
  for r in 2 5 1 6; do
    echo -n "ray defer $after: "
    timeout 120 python scripts/frame_part.py --world 8 --rank $r --steps 20 --lanes 3 --ray-defer $after | tail -1
  done
  echo -n "ray defer $after: "
  timeout 120 python scripts/frame_part.py --world 4 --rank 2 --steps 20 --lanes 3 --ray-defer $after | tail -1
done 2>&1 | tee gpurun_out/p_ab_ray_defer.txt
