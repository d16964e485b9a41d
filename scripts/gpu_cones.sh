#!/bin/bash
# normal cones: full parity suite, then A/B of the frame time and the per-bounce times
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q 2>&1 | tail -3
for v in 0 1; do
  make -C cudaraytracer_b200/csrc clean > /dev/null
  make -C cudaraytracer_b200/csrc -j16 EXTRA="-DCRT_NORMAL_CONES=$v" > /dev/null 2>&1 || { echo "build failed"; continue; }
  echo "== CRT_NORMAL_CONES=$v"
  python scripts/tune_sweep.py one 2>/dev/null | tail -1
  CRT_TRACE_TIMING=1 python scripts/quick_perf.py c3 2 > gpurun_out/cones_$v.log 2>&1
  grep -E "^(primary|incoherent)" gpurun_out/cones_$v.log
  grep -E "lane 0 bounce [0-3]:" gpurun_out/cones_$v.log | head -4
  python scripts/run_configs.py c1 c2 c3 c4 2>/dev/null | python -c "
import sys,json
for l in sys.stdin:
    d=json.loads(l); print(d['config'], round(d['ms_per_frame'],3), 'nodes/ray', round(d['nodes_per_ray'],2), 'tris/ray', round(d['tris_per_ray'],2))"
done
