#!/bin/bash
# round 2, run E (1 GPU): full parity suite, a rank's share of the frame (timeline + ncu launch list), C4 invariance at full size
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/e_pytest.log 2>&1
echo "pytest rc $?" >> gpurun_out/e_pytest.log
tail -4 gpurun_out/e_pytest.log
for w in 1 2 4 8; do python scripts/frame_part.py --world $w --steps 10; done > gpurun_out/e_parts.txt 2>&1
python scripts/frame_part.py --world 8 --lanes 1 --steps 10 >> gpurun_out/e_parts.txt 2>&1
python scripts/frame_part.py --world 8 --lanes 2 --wide-below 0 --steps 10 >> gpurun_out/e_parts.txt 2>&1
python scripts/frame_part.py --world 8 --lanes 2 --wide-below 262144 --steps 10 >> gpurun_out/e_parts.txt 2>&1
cat gpurun_out/e_parts.txt
CRT_TRACE_TIMING=1 python scripts/frame_part.py --world 8 --lanes 1 --steps 1 --timing 2> gpurun_out/e_part8_bounces.txt | tail -1
tail -24 gpurun_out/e_part8_bounces.txt
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/e_part8_launches.csv python scripts/frame_part.py --world 8 --lanes 1 --steps 2 > gpurun_out/e_ncu_run.log 2>&1
echo "ncu rc $?"
timeout 1200 python scripts/c4_invariance.py > gpurun_out/e_c4_invariance.txt 2>&1
echo "c4 invariance rc $?"
cat gpurun_out/e_c4_invariance.txt
