#!/bin/bash
# what the nvidia-smi clock sampler costs the bench (development aid)
for m in smi off smi off; do
  CRT_BENCH_SAMPLER=$m python bench.py --steps 10 --warmup 3 --no-cpu-baseline 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().split('\n')[-1]); print('$m', round(d['ms_per_step'],3), round(d['e2e']['ms_per_step'],3), d['clocks'].get('samples'))"
done
