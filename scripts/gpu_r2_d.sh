#!/bin/bash
# round 2, run D (8 GPUs): multi-rank C ABI tests, strong scaling of C3 at N = 8 and 4 (+ lane / slice variants at 8), C4 tiles over 8
mkdir -p gpurun_out
nvidia-smi -L > gpurun_out/d_gpus.txt 2>&1
timeout 900 python -m pytest tests -m gpu -q -k "ranks" > gpurun_out/d_pytest.log 2>&1
echo "pytest rc $?" >> gpurun_out/d_pytest.log
tail -4 gpurun_out/d_pytest.log
run() { # name nproc args...
  name=$1; np=$2; shift 2
  timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $np --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus $np "$@" > gpurun_out/d_$name.json 2> gpurun_out/d_$name.err
  echo "$name rc $?"
}
run c3_n8 8 --steps 20 --warmup 3
run c3_n4 4 --steps 20 --warmup 3
run c3_n8_lanes1 8 --steps 20 --warmup 3 --lanes 1
run c3_n8_slice500k 8 --steps 20 --warmup 3 --slice-min 500000
run c3_n8_slice1m 8 --steps 20 --warmup 3 --slice-min 1000000
run c3_n8_weak 8 --steps 20 --warmup 3 --scaling weak
run c4_n8 8 --steps 3 --warmup 3 --workload c4
python - <<'PY'
import json, glob
for f in sorted(glob.glob("gpurun_out/d_c*.json")):
    try:
        l = json.loads(open(f).read().strip().splitlines()[-1])
        print(f.split("/")[-1], "value", round(l["value"]), "ms", round(l["ms_per_step"], 3), "e2e ms", round(l["e2e"]["ms_per_step"], 3), l["scaling"], "lanes", l["config"]["lanes"], "parity", (l.get("parity_n") or {}).get("pixels_differing"), "launches", l["gpu_launches"])
    except Exception as e:
        print(f, "no line", e)
PY
