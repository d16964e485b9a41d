#!/bin/bash
# One full ncu capture of k_trace_queue (bounces 0..2 of a warm frame) after a plain run of the same command.
mkdir -p gpurun_out
python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/plain2.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:k_trace_queue -s 18 -c 3 -f -o gpurun_out/prof_trace python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_full.log 2>&1
tail -2 gpurun_out/ncu_full.log
