#!/bin/bash
# A/B of a compile-time switch on one box: scripts/gpu_ab.sh MACRO v1 v2 ... ; the last value stays built.
# Per value: parity subset, C3 frame time + per-bounce trace times, incoherent trace rate.
mkdir -p gpurun_out
M=$1; shift
for v in "$@"; do
  make -C cudaraytracer_b200/csrc clean > /dev/null
  make -C cudaraytracer_b200/csrc -j16 EXTRA="-D$M=$v" > /dev/null 2>&1 || { echo "build failed for $M=$v"; continue; }
  echo "== $M=$v: $(cuobjdump -res-usage cudaraytracer_b200/csrc/build/render.o 2>/dev/null | grep -A1 'k_trace_queueILb0ELb0' | grep -o 'REG:[0-9]* STACK:[0-9]* SHARED:[0-9]*')"
  python -m pytest tests -m gpu -x -q -k "brute or oracle or golden or tiny or tie or mixed or wide or oversize" 2>&1 | tail -2
  CRT_TRACE_TIMING=1 python scripts/quick_perf.py c3 2 > gpurun_out/ab_$v.log 2>&1
  grep -E "^(primary|incoherent|render)" gpurun_out/ab_$v.log
  grep -E "lane 0 bounce [0-3]:" gpurun_out/ab_$v.log | head -4
  python scripts/tune_sweep.py one 2>/dev/null | tail -1
done
