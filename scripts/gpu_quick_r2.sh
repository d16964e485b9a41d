#!/bin/bash
# slices in flight at N = 4 and N = 2: every rank's share on one GPU with at most 3 and at most 4 slices
mkdir -p gpurun_out
for lanes in 3 4; do
  for ((r = 0; r < 4; r++)); do
    echo -n "max lanes $lanes: "; timeout 120 python scripts/frame_part.py --world 4 --rank $r --steps 20 --lanes $lanes | tail -1
  done
  for ((r = 0; r < 2; r++)); do
    echo -n "max lanes $lanes: "; timeout 120 python scripts/frame_part.py --world 2 --rank $r --steps 20 --lanes $lanes | tail -1
  done
done 2>&1 | tee gpurun_out/r_ab_lanes_n4_n2.txt
