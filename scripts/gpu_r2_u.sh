#!/bin/bash
# round 2, run U (one GPU): run-time knobs after the k_shade change (slices, tile groups on C3, hand-over threshold), on the
# whole C3 frame and on a rank's share of 8; C4 with the PLOC builder on top of the tile groups
mkdir -p gpurun_out
{
timeout 300 python scripts/c3_sweep.py --world 1 --steps 30
timeout 300 python scripts/c3_sweep.py --world 8 --rank 2 --steps 40
timeout 300 python scripts/c3_sweep.py --world 8 --rank 1 --steps 40
} 2>&1 | tee gpurun_out/u_c3_sweep.txt
timeout 600 python scripts/c4_batch_ab.py --steps 3 --settings 64,0 --builder 1 2>&1 | tee gpurun_out/u_c4_ploc.txt
