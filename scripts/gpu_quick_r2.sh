#!/bin/bash
# why is one rank of 8 ~25 % slower with the same ray count? ranks 5 (slow) and 1 (fast), one slice: longest packet / ray per bounce
mkdir -p gpurun_out
for r in 5 1; do
  CRT_TRACE_TIMING=1 timeout 120 python scripts/frame_part.py --world 8 --rank $r --lanes 1 --steps 1 --timing 2>&1 | grep "nodes/ray\|bounce [012]" | head -8
done
