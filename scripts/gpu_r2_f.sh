#!/bin/bash
# round 2, run F (1 GPU): full parity suite with the new tests, C4 invariance at 64 spp after the slack change, default bench line
mkdir -p gpurun_out
timeout 2400 python -m pytest tests -m gpu -q > gpurun_out/f_pytest.log 2>&1
echo "pytest rc $?" >> gpurun_out/f_pytest.log
tail -25 gpurun_out/f_pytest.log
timeout 1200 python scripts/c4_invariance.py > gpurun_out/f_c4_invariance.txt 2>&1
echo "c4 invariance rc $?"
grep -A3 "differ" gpurun_out/f_c4_invariance.txt | head -40
timeout 900 python bench.py --steps 10 --warmup 3 > gpurun_out/f_bench.json 2> gpurun_out/f_bench.err
echo "bench rc $?"
python - <<'PY'
import json
l = json.loads(open("gpurun_out/f_bench.json").read().strip().splitlines()[-1])
print("value", round(l["value"]), "ms", round(l["ms_per_step"], 3), "e2e ms", round(l["e2e"]["ms_per_step"], 3), "frac", round(l["roofline"]["frac"], 3), "nodes/ray", round(l["roofline"]["nodes_per_ray"], 3))
print({k: (round(v["ms_per_frame"], 3) if "ms_per_frame" in v else round(v["ms_per_launch"], 3)) for k, v in l["configs"].items()})
PY
