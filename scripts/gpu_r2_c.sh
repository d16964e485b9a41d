#!/bin/bash
# round 2, run C (2 GPUs): the rotated / Cornell / multi-rank parity tests, strong-scaling bench at N = 1 and 2, crt_cli --gpus 2
mkdir -p gpurun_out
nvidia-smi -L > gpurun_out/c_gpus.txt 2>&1
timeout 900 python -m pytest tests -m gpu -q -k "rotated or cornell or ranks or communicator" > gpurun_out/c_pytest.log 2>&1
echo "pytest rc $?" >> gpurun_out/c_pytest.log
tail -15 gpurun_out/c_pytest.log
timeout 600 python bench.py --steps 10 --warmup 3 --no-cpu-baseline --no-configs > gpurun_out/c_bench_n1.json 2> gpurun_out/c_bench_n1.err
echo "bench n1 rc $?"
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 10 --warmup 3 > gpurun_out/c_bench_n2.json 2> gpurun_out/c_bench_n2.err
echo "bench n2 rc $?"
tail -3 gpurun_out/c_bench_n2.err
python - <<'PY'
import json
for n in (1, 2):
    try:
        l = json.loads(open("gpurun_out/c_bench_n%d.json" % n).read().strip().splitlines()[-1])
        print(n, "value", round(l["value"]), "ms", round(l["ms_per_step"], 3), "e2e ms", round(l["e2e"]["ms_per_step"], 3), "scaling", l["scaling"], "parity", l.get("parity_n"), "launches", l["gpu_launches"])
    except Exception as e:
        print(n, "no line", e)
PY
mkdir -p gpurun_out/cli2
timeout 300 ./cudaraytracer_b200/host/crt_cli --scene mesh --segments 400 --nx 960 --ny 540 --spp 8 --mode path --gpus 2 --frames 2 --animate --out gpurun_out/cli2/m2_ --csv gpurun_out/cli2/m2.csv > gpurun_out/c_cli2.log 2>&1
echo "cli --gpus 2 rc $?"
timeout 300 ./cudaraytracer_b200/host/crt_cli --scene mesh --segments 400 --nx 960 --ny 540 --spp 8 --mode path --frames 2 --animate --out gpurun_out/cli2/m1_ > gpurun_out/c_cli1.log 2>&1
echo "cli 1 gpu rc $?"
cmp gpurun_out/cli2/m1_0.png gpurun_out/cli2/m2_0.png && cmp gpurun_out/cli2/m1_1.png gpurun_out/cli2/m2_1.png && echo "PNG of 2 GPUs == PNG of 1 GPU"
cat gpurun_out/c_cli2.log | tail -4
