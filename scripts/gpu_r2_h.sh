#!/bin/bash
# tail hint: the new test + the tests around it, then A/B on a rank's share and the whole frame, then the bench
mkdir -p gpurun_out
timeout 600 python -m pytest tests -m gpu -x -q -k "tail_hint or wide_bounces or lanes_are or only_queue or render_frame or ranks_through or pose" > gpurun_out/h_pytest.log 2>&1
echo "pytest rc $?"; tail -5 gpurun_out/h_pytest.log
for hint in 0 1; do
  for w in 1 8; do
    for lanes in 2 4; do
      echo "hint $hint world $w lanes $lanes"
      timeout 120 python scripts/frame_part.py --world $w --steps 30 --lanes $lanes --tail-hint $hint | tail -1
    done
  done
done 2>&1 | tee gpurun_out/h_ab_tail_hint.txt
timeout 900 python bench.py --no-cpu-baseline > gpurun_out/h_bench.json 2> gpurun_out/h_bench.err
echo "bench rc $?"; cut -c1-600 gpurun_out/h_bench.json
python - <<'PY'
import json
d = json.load(open("gpurun_out/h_bench.json"))
print(d["ms_per_step"], d["e2e"]["ms_per_step"], d["roofline"]["frac"], d["gpu_launches"])
for k in d.get("other_configs", []):
    print(k)
PY
