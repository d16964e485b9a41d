#!/bin/bash
# radix tree vs PLOC topology with the normal cones (v11), C3 and C4
for b in 0 1; do echo "== CRT_BUILDER=$b"; CRT_BUILDER=$b python scripts/run_configs.py c3 c4 2>/dev/null | python -c "
import sys,json
for l in sys.stdin:
    d=json.loads(l); print(d['config'], round(d['ms_per_frame'],3), 'build', round(d['build_ms'],2), 'nodes/ray', round(d['nodes_per_ray'],2), 'tris/ray', round(d['tris_per_ray'],2))"; done
