#!/bin/bash
# round 2, run Q (8 GPUs), final code: multi-rank C ABI tests, strong scaling of C3 at N = 8 / 4 / 2, C4 over 8, crt_cli --gpus 8
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q -k "ranks" > gpurun_out/q_pytest.log 2>&1
echo "pytest rc $?"; tail -3 gpurun_out/q_pytest.log
run() { # name nproc args...
  name=$1; np=$2; shift 2
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $np --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus $np "$@" > gpurun_out/q_$name.json 2> gpurun_out/q_$name.err
  echo "$name rc $?"
}
run c3_n8 8 --steps 20 --warmup 3
run c3_n4 4 --steps 20 --warmup 3
run c3_n2 2 --steps 20 --warmup 3
run c3_n8_nodefer 8 --steps 20 --warmup 3 --packet-defer 0
run c4_n8 8 --steps 3 --warmup 3 --workload c4
( cd cudaraytracer_b200/host && timeout 300 ./crt_cli --scene mesh --gpus 8 --mode path --frames 3 --out ../../gpurun_out/q_cli8 > ../../gpurun_out/q_cli8.log 2>&1; echo "cli rc $?" )
tail -5 gpurun_out/q_cli8.log
python - <<'PY'
import json, glob
for f in sorted(glob.glob("gpurun_out/q_c*.json")):
    try:
        l = json.loads(open(f).read().strip().splitlines()[-1])
        print(f.split("/")[-1], "value", round(l["value"]), "ms", round(l["ms_per_step"], 3), "e2e ms", round(l["e2e"]["ms_per_step"], 3), l["scaling"], "lanes", l["config"]["lanes"], "parity", (l.get("parity_n") or {}).get("pixels_differing"), "launches", l["gpu_launches"])
        print("   per_rank", l["config"].get("per_rank"))
    except Exception as e:
        print(f, "no line", e)
PY
