#!/bin/bash
# does a rank's share scale worse because the tree no longer fits the caches? small mesh (fits L2) vs the 1 M-triangle one
mkdir -p gpurun_out
for seg in 200 1000; do
  for w in 1 8; do
    timeout 120 python scripts/frame_part.py --world $w --steps 20 --lanes 1 --segments $seg | tail -1
  done
done
CRT_TRACE_TIMING=1 timeout 120 python scripts/frame_part.py --world 8 --lanes 1 --steps 1 --segments 200 --timing 2>&1 | grep "lane 0 bounce [012]" | tail -3
CRT_TRACE_TIMING=1 timeout 120 python scripts/frame_part.py --world 1 --lanes 1 --steps 1 --segments 200 --timing 2>&1 | grep "lane 0 bounce [012]" | tail -3
CRT_TRACE_TIMING=1 timeout 120 python scripts/frame_part.py --world 1 --lanes 1 --steps 1 --timing 2>&1 | grep "lane 0 bounce [012]" | tail -3
