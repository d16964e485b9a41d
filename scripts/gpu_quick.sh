#!/bin/bash
# Short GPU pass: parity tests, quick perf probe, bench.
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_gpu.log
CRT_TRACE_TIMING=1 python scripts/quick_perf.py c3 > gpurun_out/quick_c3.log 2>&1
python bench.py --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench exit $?" >> gpurun_out/bench.err
tail -3 gpurun_out/pytest_gpu.log; tail -32 gpurun_out/quick_c3.log; cat gpurun_out/bench.json
