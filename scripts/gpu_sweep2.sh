#!/bin/bash
# leaf_max / refill / leaf_thresh sweep of the render path on C3 (development aid), one slice in flight so the
# per-bounce times are those of the kernels alone.
for cfg in "1 8 8" "2 8 8" "3 8 8" "4 8 8" "2 4 8" "2 12 8" "2 16 8" "2 8 4" "2 8 12" "2 8 16" "3 8 12"; do
  echo "== leaf_max/refill/leaf_thresh $cfg"
  CRT_LANES=1 CRT_TRACE_TIMING=1 python scripts/quick_perf.py c3 $cfg 2>&1 | grep -E "bounce [012]:|render path" | tail -5 | head -4 | cut -c1-110
done
