#!/bin/bash
# occupancy sweep of the per-ray traversal kernel (development aid).
mkdir -p gpurun_out
for mb in 6 8 9 10 12; do
  make -C cudaraytracer_b200/csrc clean > /dev/null
  make -C cudaraytracer_b200/csrc -j8 EXTRA="-DCRT_TRACE_MIN_BLOCKS=$mb" > /dev/null 2>&1
  echo "== min_blocks $mb: $(cuobjdump -res-usage cudaraytracer_b200/csrc/build/render.o 2>/dev/null | grep -A1 'k_trace_queueILb0ELb0' | grep -o 'REG:[0-9]*\|STACK:[0-9]*' | paste - -)"
  CRT_LANES=1 CRT_TRACE_TIMING=1 python scripts/quick_perf.py ${SCENE:-c3} 2 8 8 2>&1 | grep -E "bounce [12]:|render path" | tail -4 | head -3 | cut -c1-100
done
make -C cudaraytracer_b200/csrc clean > /dev/null
make -C cudaraytracer_b200/csrc -j8 > /dev/null 2>&1
