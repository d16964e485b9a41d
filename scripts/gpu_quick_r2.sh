#!/bin/bash
# fewer resident warps for the camera-ray launch: every warp (and the longest packet) gets a larger share of its scheduler
mkdir -p gpurun_out
for bps in 8 6 4 3; do
  export CRT_PACKET_BPS=$bps
  for cfg in "8 2" "8 6" "1 0"; do
    set -- $cfg
    echo -n "packet blocks/SM $bps: "
    timeout 120 python scripts/frame_part.py --world $1 --rank $2 --steps 20 --lanes 2 | tail -1
  done
done 2>&1 | tee gpurun_out/m_ab_packet_bps.txt
unset CRT_PACKET_BPS
for bps in 6 4; do
  export CRT_TRACE_BPS=$bps
  for cfg in "8 2" "8 6" "1 0"; do
    set -- $cfg
    echo -n "queue-kernel blocks/SM $bps: "
    timeout 120 python scripts/frame_part.py --world $1 --rank $2 --steps 20 --lanes 2 | tail -1
  done
done 2>&1 | tee -a gpurun_out/m_ab_packet_bps.txt
