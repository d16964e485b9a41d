#!/bin/bash
# parity of crt_render_frame + e2e comparison of the two end-to-end forms
mkdir -p gpurun_out
python -m pytest tests -x -q -m gpu -k "frame or lanes or tiles or golden or c1" 2>&1 | tail -5
python bench.py --steps 10 --warmup 3 --no-cpu-baseline --e2e-api calls 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().split('\n')[-1]); print('calls', d['ms_per_step'], d['e2e']['ms_per_step'], d['e2e']['value'])"
python bench.py --steps 10 --warmup 3 --no-cpu-baseline --e2e-api frame 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().split('\n')[-1]); print('frame', d['ms_per_step'], d['e2e']['ms_per_step'], d['e2e']['value'])"
