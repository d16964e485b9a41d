#!/bin/bash
# Lane-count sweep of the render path on C3 (CRT_OPT_LANES).
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_gpu.log
tail -3 gpurun_out/pytest_gpu.log
for l in 1 2 3 4; do
  echo "== lanes $l"
  CRT_LANES=$l python scripts/quick_perf.py c3 2>&1 | grep "^render"
done
CRT_LANES=2 CRT_TRACE_TIMING=1 python scripts/quick_perf.py c3 2>&1 | grep "lane" | tail -22
