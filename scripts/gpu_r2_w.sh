#!/bin/bash
# round 2, run W (one GPU): A/B of library builds under build_ab/ - k_shade prefetching the next iteration's queue entry into L2;
# 64 / 256 threads per block in the wavefront kernels (k_raygen, k_shade)
mkdir -p gpurun_out
{
for v in "" prefetch rb64 rb256 ""; do
  if [ -z "$v" ]; then unset CRT_LIB; name="current"; else export CRT_LIB=$PWD/build_ab/libcrt_$v.so; name=$v; fi
  [ -n "$v" ] && [ ! -f "$CRT_LIB" ] && continue
  echo "== $name"
  timeout 300 python scripts/frame_part.py --world 1 --lanes 3 --steps 40 | tail -1
  timeout 300 python scripts/frame_part.py --world 8 --lanes 3 --steps 40 --ranks 1,2 | tail -2
  if [ -n "$v" ]; then
    timeout 600 python -m pytest tests -m gpu -x -q -k "golden or bunny or cornell or lanes_are or tile_groups" 2>&1 | tail -1
  fi
done
unset CRT_LIB
} 2>&1 | tee gpurun_out/w_ab.txt
