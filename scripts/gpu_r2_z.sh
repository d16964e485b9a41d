#!/bin/bash
# round 2, final run (one GPU): what the driver runs at round end - GPU tests, smoke, the bench line, the reference arm
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/z_pytest.log 2>&1; echo "pytest rc $?" | tee -a gpurun_out/z_pytest.log; tail -3 gpurun_out/z_pytest.log
timeout 600 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/z_smoke.log 2>&1; echo "smoke rc $?" | tee -a gpurun_out/z_smoke.log; tail -2 gpurun_out/z_smoke.log
t0=$SECONDS; timeout 1200 python bench.py > gpurun_out/z_bench.json 2> gpurun_out/z_bench.err; echo "bench rc $? in $((SECONDS - t0)) s"
t0=$SECONDS; timeout 1200 python bench.py --impl reference > gpurun_out/z_bench_reference.json 2> gpurun_out/z_bench_reference.err; echo "reference arm rc $? in $((SECONDS - t0)) s"
python - <<'PY'
import json
d = json.loads(open("gpurun_out/z_bench.json").read().strip().splitlines()[-1])
print("value", round(d["value"]), "ms", round(d["ms_per_step"], 3), "e2e", round(d["e2e"]["value"]), round(d["e2e"]["ms_per_step"], 3), "frac", round(d["roofline"]["frac"], 3),
      "traffic", d["roofline"]["traffic"], "launches", d["gpu_launches"], "clocks", d["clocks"])
print("cpu_baseline", {k: d["cpu_baseline"][k] for k in ("value", "cores", "kind", "bvh_build_s", "rmse_gpu_vs_reference", "pixels_differing") if k in d["cpu_baseline"]})
for k, v in d["configs"].items():
    print(k, {a: (round(b, 3) if isinstance(b, float) else b) for a, b in v.items() if a in ("ms_per_frame", "ms_per_launch", "mrays_per_s", "roofline_frac")})
r = json.loads(open("gpurun_out/z_bench_reference.json").read().strip().splitlines()[-1])
print("reference arm", r["value"], r["unit"], r["ms_per_step"], r["cpu_baseline"]["cores"], r["cpu_baseline"]["sample"][:120])
PY
