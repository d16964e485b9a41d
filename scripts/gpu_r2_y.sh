#!/bin/bash
# round 2, run Y (eight GPUs): C4 split by tiles over 8 ranks with the tile groups inside each rank's share and the pose as
# one translation per instance (the C3 scaling line is the driver's own run at round end)
mkdir -p gpurun_out
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29513 bench.py --gpus 8 --workload c4 --steps 5 --warmup 3 > gpurun_out/y_c4_n8.json 2> gpurun_out/y_c4_n8.err; echo "c4 n8 rc $?"
python - <<'PY'
import json
f = "gpurun_out/y_c4_n8.json"
try:
    d = json.loads([l for l in open(f).read().strip().splitlines() if l.startswith("{")][-1])
    print(f, "ms", round(d["ms_per_step"], 3), "value", round(d["value"]), "e2e ms", round(d["e2e"]["ms_per_step"], 3), "parity", d.get("parity_n", {}).get("pixels_differing"), d["config"]["per_rank"])
except Exception as e:
    print(f, "no line:", e)
PY
