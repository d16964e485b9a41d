#!/bin/bash
# A/B of a compile-time switch by C3 frame time: scripts/gpu_ab_frame.sh MACRO v1 v2 ...; the last value stays built
mkdir -p gpurun_out
M=$1; shift
for v in "$@"; do
  make -C cudaraytracer_b200/csrc clean > /dev/null
  make -C cudaraytracer_b200/csrc -j16 EXTRA="-D$M=$v" > /dev/null 2>&1 || { echo "build failed for $M=$v"; continue; }
  echo "== $M=$v"
  python -m pytest tests -m gpu -x -q -k "golden or bit_identical or lanes or tiles" 2>&1 | tail -1
  python scripts/tune_sweep.py one 2>/dev/null | tail -1
  python scripts/tune_sweep.py one 2>/dev/null | tail -1
done
