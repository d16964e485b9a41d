#!/bin/bash
# stream priority modes of crt_create (development aid): 0 = none, 1 = descending lane priorities, 2 = only upload + refit raised
for p in ${@:-0 2 0 2}; do echo "== CRT_LANE_PRIO=$p"; CRT_LANE_PRIO=$p python scripts/e2e_probe.py 2>&1 | head -6; done
