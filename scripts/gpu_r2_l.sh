#!/bin/bash
# longest camera-ray packets first (CRT_OPT_PACKET_SCHED): parity, then every rank's share of 8 on one GPU and the whole frame, on and off
mkdir -p gpurun_out
timeout 600 python -m pytest tests -m gpu -x -q -k "packet_schedule or tail_hint or wide_bounces or lanes_are or packet" > gpurun_out/l_pytest.log 2>&1
echo "pytest rc $?"; tail -5 gpurun_out/l_pytest.log
for sched in 0 128; do
  for ((r = 0; r < 8; r++)); do
    echo -n "sched $sched: "
    timeout 120 python scripts/frame_part.py --world 8 --rank $r --steps 20 --lanes 2 --packet-sched $sched | tail -1
  done
  for lanes in 2 4; do
    echo -n "sched $sched: "
    timeout 120 python scripts/frame_part.py --world 1 --steps 30 --lanes $lanes --packet-sched $sched | tail -1
  done
done 2>&1 | tee gpurun_out/l_ab_packet_sched.txt
for sched in 64 256; do
  for r in 5 1; do
    echo -n "sched $sched: "
    timeout 120 python scripts/frame_part.py --world 8 --rank $r --steps 20 --lanes 2 --packet-sched $sched | tail -1
  done
done 2>&1 | tee -a gpurun_out/l_ab_packet_sched.txt
