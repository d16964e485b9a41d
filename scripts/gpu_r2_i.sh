#!/bin/bash
# round 2, run I (8 GPUs): strong scaling of C3 with the tile gather against the reduce, tail hint on, C4 over 8
mkdir -p gpurun_out
run() { # name nproc args...
  name=$1; np=$2; shift 2
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $np --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus $np "$@" > gpurun_out/i_$name.json 2> gpurun_out/i_$name.err
  echo "$name rc $?"
}
run c3_n8 8 --steps 20 --warmup 3
run c3_n8_reduce 8 --steps 20 --warmup 3 --combine reduce
run c3_n4 4 --steps 20 --warmup 3
run c3_n2 2 --steps 20 --warmup 3
run c4_n8 8 --steps 3 --warmup 3 --workload c4
python - <<'PY'
import json, glob
for f in sorted(glob.glob("gpurun_out/i_c*.json")):
    try:
        l = json.loads(open(f).read().strip().splitlines()[-1])
        print(f.split("/")[-1], "value", round(l["value"]), "ms", round(l["ms_per_step"], 3), "e2e ms", round(l["e2e"]["ms_per_step"], 3), l["scaling"], "lanes", l["config"]["lanes"], "parity", (l.get("parity_n") or {}).get("pixels_differing"), "launches", l["gpu_launches"])
        print("   per_rank", l["config"].get("per_rank"))
    except Exception as e:
        print(f, "no line", e)
PY
