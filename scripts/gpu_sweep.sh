#!/bin/bash
# occupancy sweep of the traversal kernels (development aid).
mkdir -p gpurun_out
for mb in 5 6 7; do
  make -C cudaraytracer_b200/csrc clean > /dev/null
  make -C cudaraytracer_b200/csrc -j8 EXTRA="-DCRT_TRACE_MIN_BLOCKS=$mb -Xptxas -v" 2>&1 | grep -A2 "k_trace_queueILb0ELb0\|k_trace_packetILb0ELb0" | grep registers
  for cfg in "2 8 8" "2 12 8"; do
    echo "== min_blocks $mb leaf_max/refill/leaf_thresh $cfg"
    CRT_TRACE_TIMING=1 python scripts/quick_perf.py ${SCENE:-c3} $cfg 2>&1 | grep -E "bounce [01]:|render path" | tail -4 | head -3
  done
done
make -C cudaraytracer_b200/csrc clean > /dev/null
make -C cudaraytracer_b200/csrc -j8 > /dev/null 2>&1
