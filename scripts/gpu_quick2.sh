#!/bin/bash
# parity tests + every config
mkdir -p gpurun_out
timeout 600 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_gpu.log
tail -12 gpurun_out/pytest_gpu.log
timeout 900 python scripts/run_configs.py c1 c2 c3 c4 c5 > gpurun_out/configs.jsonl 2> gpurun_out/configs.err
python - <<'PY'
import json
for l in open('gpurun_out/configs.jsonl'):
    d=json.loads(l)
    print('%-30s prims %9d ms %9.3f Mrays/s %8.1f nodes/ray %6.2f tris/ray %5.2f alg GB/s %7.1f' % (d['config'], d['prims'], d.get('ms_per_frame', d.get('ms', 0)), d['mrays_s'], d['nodes_per_ray'], d['tris_per_ray'], d['alg_gbs']))
PY
tail -3 gpurun_out/configs.err
