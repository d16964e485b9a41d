#!/bin/bash
# full ncu capture of k_raygen and the first two k_shade launches of one warm frame (one slice in flight)
mkdir -p gpurun_out
python bench.py --steps 2 --warmup 3 --no-cpu-baseline --lanes 1 > gpurun_out/plain3.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:"k_shade|k_raygen" -s 20 -c 3 -o gpurun_out/prof_shade -f python bench.py --steps 2 --warmup 3 --no-cpu-baseline --lanes 1 > gpurun_out/ncu_shade.log 2>&1
tail -3 gpurun_out/ncu_shade.log | cut -c1-200
