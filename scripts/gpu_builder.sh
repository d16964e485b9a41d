#!/bin/bash
# LBVH vs PLOC topology (CRT_BUILDER / CRT_PLOC_RADIUS environment overrides) on every config.
mkdir -p gpurun_out
for cfg in "0 16" "1 8" "1 16" "1 32"; do
  set -- $cfg
  echo "== builder $1 radius $2"
  CRT_BUILDER=$1 CRT_PLOC_RADIUS=$2 timeout 600 python scripts/run_configs.py c1 c2 c3 c4 2>/dev/null | python -c "
import sys, json
for l in sys.stdin:
    d = json.loads(l)
    print('%-30s ms %8.3f  nodes/ray %6.2f tris/ray %5.2f sph/ray %5.2f build %6.2f' % (d['config'], d.get('ms_per_frame', d.get('ms', 0)), d['nodes_per_ray'], d['tris_per_ray'], d.get('spheres_per_ray', 0), d['build_ms']))
"
done
