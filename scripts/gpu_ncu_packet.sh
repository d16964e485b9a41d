#!/bin/bash
mkdir -p gpurun_out
python bench.py --steps 2 --warmup 3 --no-cpu-baseline --lanes 1 > gpurun_out/plain4.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:k_trace_packet -s 2 -c 1 -o gpurun_out/prof_packet -f python bench.py --steps 2 --warmup 3 --no-cpu-baseline --lanes 1 > gpurun_out/ncu_packet.log 2>&1
tail -2 gpurun_out/ncu_packet.log | cut -c1-200
