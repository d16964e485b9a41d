#!/bin/bash
# long camera-ray packets handed on to a warp-per-ray launch (CRT_OPT_PACKET_DEFER, slices <= 4 M samples): full GPU suite, then every rank's share
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/p_pytest.log 2>&1
echo "pytest rc $?"; tail -4 gpurun_out/p_pytest.log
for after in 0 192; do
  for ((r = 0; r < 8; r++)); do
    echo -n "defer $after: "
    timeout 120 python scripts/frame_part.py --world 8 --rank $r --steps 20 --lanes 3 --packet-defer $after | tail -1
  done
  for cfg in "4 2 3" "4 0 3" "2 0 3" "1 0 3"; do
    set -- $cfg
    echo -n "defer $after: "
    timeout 120 python scripts/frame_part.py --world $1 --rank $2 --steps 20 --lanes $3 --packet-defer $after | tail -1
  done
done 2>&1 | tee gpurun_out/p_ab_packet_defer2.txt
