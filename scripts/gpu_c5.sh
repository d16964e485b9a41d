#!/bin/bash
timeout 600 python scripts/run_configs.py c1 c5 2>/dev/null | python -c "
import sys, json
for l in sys.stdin:
    d = json.loads(l)
    print('%-30s prims %9d ms %8.3f  nodes/ray %6.2f tris/ray %5.2f' % (d['config'], d['prims'], d.get('ms_per_frame', d.get('ms', 0)), d['nodes_per_ray'], d['tris_per_ray']))
"
CRT_LANES=1 CRT_TRACE_TIMING=1 python scripts/quick_perf.py c3 2>&1 | grep -E "bounce [01]:|render" | tail -6
