#!/bin/bash
# drain mode of the per-ray launches: parity, then a sweep of CRT_OPT_DRAIN_AFTER on the slow and a fast rank of 8 and the whole frame
mkdir -p gpurun_out
timeout 600 python -m pytest tests -m gpu -x -q -k "drain or tail_hint or wide_bounces or lanes_are" > gpurun_out/k_pytest.log 2>&1
echo "pytest rc $?"; tail -5 gpurun_out/k_pytest.log
for after in 0 16 32 64 128; do
  for cfg in "8 5" "8 1" "1 0"; do
    set -- $cfg
    echo -n "drain_after $after: "
    timeout 120 python scripts/frame_part.py --world $1 --rank $2 --steps 20 --lanes 2 --drain-after $after | tail -1
  done
done 2>&1 | tee gpurun_out/k_ab_drain.txt
for after in 0 32; do
  CRT_TRACE_TIMING=1 timeout 120 python scripts/frame_part.py --world 8 --rank 5 --lanes 1 --steps 1 --timing --drain-after $after 2>&1 | grep "lane 0 bounce [012]" | tail -3
done 2>&1 | tee -a gpurun_out/k_ab_drain.txt
