#!/bin/bash
# One GPU-box pass: parity tests, smoke, bench, launch list, one full ncu capture.
set -x
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_gpu.log
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke exit $?" >> gpurun_out/smoke.log
python bench.py > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench exit $?" >> gpurun_out/bench.err
CRT_TRACE_TIMING=1 python scripts/quick_perf.py c3 > gpurun_out/quick_c3.log 2>&1
python scripts/run_configs.py c1 c2 c3 c4 c5 > gpurun_out/configs.jsonl 2> gpurun_out/configs.err
python bench.py --steps 2 --warmup 3 --no-cpu-baseline --lanes 1 > gpurun_out/plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline --lanes 1 > gpurun_out/ncu_launch.log 2>&1
python bench.py --steps 2 --warmup 3 --no-cpu-baseline --lanes 1 > gpurun_out/plain2.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:k_trace_ -s 18 -c 3 -o gpurun_out/prof_trace python bench.py --steps 2 --warmup 3 --no-cpu-baseline --lanes 1 > gpurun_out/ncu_full.log 2>&1
ls -la gpurun_out
