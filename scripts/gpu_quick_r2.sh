#!/bin/bash
# register budgets: queue kernel at 72 registers (min blocks 5), packet kernel at 72 (min blocks 6), against the default (64 / 62)
mkdir -p gpurun_out
for lib in default t5 p6; do
  if [ $lib = default ]; then unset CRT_LIB; else export CRT_LIB=$PWD/cudaraytracer_b200/libcrt_$lib.so; fi
  for lanes in 2 4; do
    echo -n "lib $lib: "
    timeout 120 python scripts/frame_part.py --world 1 --steps 30 --lanes $lanes | tail -1
  done
done 2>&1 | tee gpurun_out/o_ab_registers.txt
unset CRT_LIB
for lanes in 2 3 4; do
  timeout 300 python bench.py --no-cpu-baseline --no-configs --lanes $lanes > gpurun_out/o_bench_lanes$lanes.json 2>/dev/null
  python - <<PY
import json
d = json.load(open("gpurun_out/o_bench_lanes$lanes.json"))
print("bench lanes $lanes: value ms", round(d["ms_per_step"], 3), "e2e ms", round(d["e2e"]["ms_per_step"], 3), "frac", round(d["roofline"]["frac"], 3), "traffic", d["roofline"]["traffic"])
PY
done 2>&1 | tee -a gpurun_out/o_ab_registers.txt
