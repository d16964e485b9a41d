#!/bin/bash
# occupancy / leaf-size / leaf-threshold sweep of the traversal kernels (development aid).
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q 2>&1 | tail -3
for mb in 9 8; do
  make -C cudaraytracer_b200/csrc clean > /dev/null
  make -C cudaraytracer_b200/csrc -j8 EXTRA="-DCRT_TRACE_MIN_BLOCKS=$mb" > /dev/null 2>&1
  for cfg in "2 8 8" "1 8 8" "4 8 8" "2 8 4" "2 8 12" "2 4 8" "2 12 8"; do
    echo "== min_blocks $mb leaf_max/refill/leaf_thresh $cfg"
    python scripts/quick_perf.py ${SCENE:-c3} $cfg 2>&1 | grep -E "primary|incoherent|render path"
  done
done
make -C cudaraytracer_b200/csrc clean > /dev/null
make -C cudaraytracer_b200/csrc -j8 > /dev/null 2>&1
