#!/bin/bash
# round 2, run ZZ (one GPU, the round's last GPU seconds): the build with CRT_QUEUE_LOCAL_DIR=1 (build_ab/libcrt_localdir.so) -
# the whole GPU test suite on it, then its C3 frame and its roofline pass beside the current default
mkdir -p gpurun_out
export CRT_LIB=$PWD/build_ab/libcrt_localdir.so
timeout 100 python -m pytest tests -m gpu -x -q > gpurun_out/zz_pytest.log 2>&1; echo "pytest (localdir) rc $?" | tee -a gpurun_out/zz_pytest.log; tail -3 gpurun_out/zz_pytest.log
{
echo "== localdir"; timeout 40 python scripts/frame_part.py --world 1 --lanes 3 --steps 30 | tail -1
unset CRT_LIB
echo "== current"; timeout 40 python scripts/frame_part.py --world 1 --lanes 3 --steps 30 | tail -1
} 2>&1 | tee gpurun_out/zz_ab.txt
export CRT_LIB=$PWD/build_ab/libcrt_localdir.so
timeout 60 python bench.py --no-cpu-baseline --no-configs > gpurun_out/zz_bench_localdir.json 2> gpurun_out/zz_bench_localdir.err
python - <<'PY'
import json
try:
    d = json.loads(open("gpurun_out/zz_bench_localdir.json").read().strip().splitlines()[-1])
    print("localdir bench: value", round(d["value"]), "ms", round(d["ms_per_step"], 3), "e2e ms", round(d["e2e"]["ms_per_step"], 3), "frac", round(d["roofline"]["frac"], 4), "avg launch ms", round(d["roofline"]["avg_launch_ms"], 4))
except Exception as e:
    print("no bench line:", e)
PY
