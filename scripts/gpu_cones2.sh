#!/bin/bash
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q 2>&1 | tail -2
python scripts/tune_sweep.py one 2>/dev/null | tail -1
python scripts/run_configs.py c1 c3 c4 2>/dev/null | python -c "
import sys,json
for l in sys.stdin:
    d=json.loads(l); print(d['config'], round(d['ms_per_frame'],3), 'nodes/ray', round(d['nodes_per_ray'],2), 'tris/ray', round(d['tris_per_ray'],2))"
