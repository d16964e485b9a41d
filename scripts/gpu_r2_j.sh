#!/bin/bash
# tile numbering with rotated rows: every rank's share of 8 (and 4) on ONE GPU, old numbering (CRT_TILE_ROT=0) against the new
mkdir -p gpurun_out
timeout 600 python -m pytest tests -m gpu -x -q -k "rank or lanes_are or tiles or part or comm or tail_hint" > gpurun_out/j_pytest.log 2>&1
echo "pytest rc $?"; tail -3 gpurun_out/j_pytest.log
for rot in 0 default; do
  for w in 8 4; do
    for ((r = 0; r < w; r++)); do
      if [ $rot = 0 ]; then export CRT_TILE_ROT=0; else unset CRT_TILE_ROT; fi
      echo -n "rot $rot: "
      timeout 120 python scripts/frame_part.py --world $w --rank $r --steps 20 --lanes 2 | tail -1
    done
  done
done 2>&1 | tee gpurun_out/j_ab_tile_rotation.txt
unset CRT_TILE_ROT
for lanes in 2 4; do timeout 120 python scripts/frame_part.py --world 1 --steps 30 --lanes $lanes | tail -1; done | tee -a gpurun_out/j_ab_tile_rotation.txt
CRT_TILE_ROT=0 timeout 120 python scripts/frame_part.py --world 1 --steps 30 --lanes 4 | tail -1 | tee -a gpurun_out/j_ab_tile_rotation.txt
