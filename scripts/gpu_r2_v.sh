#!/bin/bash
# round 2, final run V (one GPU): what the driver runs at round end (GPU tests, smoke, the bench line, the reference arm),
# then the ncu evidence of this code: launch list of the bench command and a full capture of one warm C3 frame.
# Every profiled command first exits 0 without ncu. scripts/ncu_summary.py gpurun_out/v r02_v turns the exports into profiles/.
mkdir -p gpurun_out/v
O=gpurun_out/v
t0=$SECONDS
timeout 1200 python -m pytest tests -m gpu -x -q > $O/pytest.log 2>&1; echo "pytest rc $?" | tee -a $O/pytest.log; tail -3 $O/pytest.log
timeout 600 python -c "import __graft_entry__ as g; g.smoke()" > $O/smoke.log 2>&1; echo "smoke rc $?" | tee -a $O/smoke.log; tail -2 $O/smoke.log
timeout 1200 python bench.py > $O/bench.json 2> $O/bench.err; echo "bench rc $? [$((SECONDS - t0)) s]"
timeout 1200 python bench.py --impl reference > $O/bench_reference.json 2> $O/bench_reference.err; echo "reference arm rc $? [$((SECONDS - t0)) s]"
python - <<'PY'
import json
d = json.loads(open("gpurun_out/v/bench.json").read().strip().splitlines()[-1])
print("value", round(d["value"]), "ms", round(d["ms_per_step"], 3), "e2e", round(d["e2e"]["value"]), round(d["e2e"]["ms_per_step"], 3), "frac", round(d["roofline"]["frac"], 3),
      "traffic", d["roofline"]["traffic"], "launches", d["gpu_launches"], "clocks", d["clocks"])
print("cpu_baseline", {k: d["cpu_baseline"][k] for k in ("value", "cores", "kind", "bvh_build_s", "rmse_gpu_vs_reference", "pixels_differing") if k in d["cpu_baseline"]})
for k, v in d["configs"].items():
    print(k, {a: (round(b, 3) if isinstance(b, float) else b) for a, b in v.items() if a in ("ms_per_frame", "ms_per_launch", "mrays_per_s", "roofline_frac")})
r = json.loads(open("gpurun_out/v/bench_reference.json").read().strip().splitlines()[-1])
print("reference arm", r["value"], r["unit"], r["ms_per_step"], r["cpu_baseline"]["cores"], r["cpu_baseline"]["sample"][:120])
PY
NCU="ncu --clock-control none"
timeout 600 python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-configs > $O/n_plain_bench.json 2> $O/n_plain_bench.err && \
timeout 900 $NCU --metrics gpu__time_duration.sum -c 700 --csv --log-file $O/n_launches.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-configs > $O/n_ncu_launch.log 2>&1
echo "launch list rc $? [$((SECONDS - t0)) s]"
timeout 300 python scripts/ncu_frame.py > $O/n_frame_plain.log 2>&1 && cat $O/n_frame_plain.log && \
timeout 1500 $NCU --profile-from-start off --set full --import-source on -f -o $O/n_frame python scripts/ncu_frame.py > $O/n_frame_ncu.log 2>&1
echo "frame capture rc $? [$((SECONDS - t0)) s]"
ncu -i $O/n_frame.ncu-rep --page raw --csv > $O/n_frame_raw.csv 2> /dev/null
rm -f $O/n_frame.ncu-rep
ls -la $O
