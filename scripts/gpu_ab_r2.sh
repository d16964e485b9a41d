#!/bin/bash
# A/B of library builds under build_ab/ (CRT_LIB selects the .so): C3 frame, a rank's share, work counters
mkdir -p gpurun_out
for v in "" noD3 old pad21 slack21 $AB_EXTRA; do
  if [ -z "$v" ]; then unset CRT_LIB; name=current; else export CRT_LIB=$PWD/build_ab/libcrt_$v.so; name=$v; fi
  [ -n "$v" ] && [ ! -f "$CRT_LIB" ] && continue
  echo "== $name"
  python scripts/frame_part.py --world 1 --steps 20
  python scripts/frame_part.py --world 1 --steps 20 --lanes 1 | tail -1
done 2>&1 | tee gpurun_out/ab.txt
