#!/bin/bash
# quick A/B: parity subset + C3 frame time + a rank's share
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q -k "bunny or spheres or c3 or rotated or fuzz or golden" > gpurun_out/q_pytest.log 2>&1
echo "pytest rc $?"; tail -2 gpurun_out/q_pytest.log
for w in 1 8; do python scripts/frame_part.py --world $w --steps 20; done
python scripts/frame_part.py --world 1 --workload c4 --steps 3
