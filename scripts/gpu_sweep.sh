#!/bin/bash
# leaf-size / refill / leaf-threshold sweep of the traversal kernels (development aid).
mkdir -p gpurun_out
for cfg in "2 8 4" "2 8 6" "2 8 8" "2 8 12" "3 8 4" "3 8 6" "3 8 8" "3 8 12" "4 8 4" "4 8 6" "4 8 10" "4 6 8" "4 10 8" "2 6 8" "1 8 8"; do
  echo "== leaf_max/refill/leaf_thresh $cfg"
  python scripts/quick_perf.py ${SCENE:-c3} $cfg 2>&1 | grep -E "primary|incoherent|render path"
done
