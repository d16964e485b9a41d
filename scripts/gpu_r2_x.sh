#!/bin/bash
# round 2, run X (two GPUs): the N > 1 path after the tile groups and the per-group tail hints - the C-ABI-only rank test,
# the strong-scaling bench line with its in-run parity check, and C4 over two ranks (tile groups inside each rank's share)
mkdir -p gpurun_out
timeout 600 python -m pytest tests -m gpu -x -q -k "ranks_through or communicator" 2>&1 | tail -3
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 20 --warmup 3 > gpurun_out/x_c3_n2.json 2> gpurun_out/x_c3_n2.err; echo "c3 n2 rc $?"
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus 2 --workload c4 --steps 3 --warmup 3 > gpurun_out/x_c4_n2.json 2> gpurun_out/x_c4_n2.err; echo "c4 n2 rc $?"
python - <<'PY'
import json
for f in ("gpurun_out/x_c3_n2.json", "gpurun_out/x_c4_n2.json"):
    try:
        d = json.loads([l for l in open(f).read().strip().splitlines() if l.startswith("{")][-1])
        print(f, "ms", round(d["ms_per_step"], 3), "value", round(d["value"]), "e2e ms", round(d["e2e"]["ms_per_step"], 3), "parity", d.get("parity_n", {}).get("pixels_differing"), d["config"]["per_rank"])
    except Exception as e:
        print(f, "no line:", e)
PY
