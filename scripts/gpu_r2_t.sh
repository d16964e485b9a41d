#!/bin/bash
# round 2, run T (one GPU): k_shade at 48 registers as the default, a tail hint per tile group, and an A/B of library
# builds under build_ab/ (k_shade at 56 = the old default and at 40 registers; 8 / 16 shared-memory levels of the per-ray stack)
mkdir -p gpurun_out
t0=$SECONDS
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/t_pytest.log 2>&1; echo "pytest rc $?" | tee -a gpurun_out/t_pytest.log; tail -3 gpurun_out/t_pytest.log
echo "[$((SECONDS - t0)) s]"
{
for v in "" shade9 shade12 stack8 stack16; do
  if [ -z "$v" ]; then unset CRT_LIB; name="current (k_shade at 48 registers)"; else export CRT_LIB=$PWD/build_ab/libcrt_$v.so; name=$v; fi
  [ -n "$v" ] && [ ! -f "$CRT_LIB" ] && continue
  echo "== $name"
  timeout 300 python scripts/frame_part.py --world 1 --lanes 3 --steps 30
  timeout 300 python scripts/frame_part.py --world 8 --lanes 3 --steps 30 --ranks 0,1,2 | tail -3
  if [ -n "$v" ] && [ "$v" != shade9 ]; then
    timeout 600 python -m pytest tests -m gpu -x -q -k "golden or bunny or packet or cornell or lanes_are or c3" 2>&1 | tail -2
  fi
done
unset CRT_LIB
} 2>&1 | tee gpurun_out/t_ab.txt
echo "[$((SECONDS - t0)) s]"
timeout 900 python scripts/c4_batch_ab.py --steps 3 --settings 0,64,32 2>&1 | tee gpurun_out/t_c4_batch_ab.txt
echo "[$((SECONDS - t0)) s]"
timeout 900 python bench.py --workload c4 --steps 3 --warmup 3 > gpurun_out/t_bench_c4.json 2> gpurun_out/t_bench_c4.err; echo "bench c4 rc $?"
python - <<'PY'
import json
try:
    d = json.loads(open("gpurun_out/t_bench_c4.json").read().strip().splitlines()[-1])
    print("c4: value", round(d["value"]), "ms", round(d["ms_per_step"], 2), "e2e ms", round(d["e2e"]["ms_per_step"], 2), "h2d", d["e2e"]["h2d_bytes_per_step"], "frac", round(d["roofline"]["frac"], 3))
except Exception as e:
    print("no c4 line:", e)
PY
echo "[$((SECONDS - t0)) s]"
