#!/bin/bash
# round 2, run N (one GPU): the ncu evidence of the final code. Every profiled command first exits 0 without ncu.
mkdir -p gpurun_out
NCU="ncu --clock-control none"
# 1. launch list of the bench command
timeout 600 python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-configs > gpurun_out/n_plain_bench.json 2> gpurun_out/n_plain_bench.err && \
timeout 900 $NCU --metrics gpu__time_duration.sum -c 700 --csv --log-file gpurun_out/n_launches.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-configs > gpurun_out/n_ncu_launch.log 2>&1
echo "launch list rc $?"
# 2. full capture of one warm frame, one slice in flight, every kernel of the frame
timeout 300 python scripts/ncu_frame.py > gpurun_out/n_frame_plain.log 2>&1 && cat gpurun_out/n_frame_plain.log && \
timeout 1500 $NCU --profile-from-start off --set full --import-source on -f -o gpurun_out/n_frame python scripts/ncu_frame.py > gpurun_out/n_frame_ncu.log 2>&1
echo "frame capture rc $?"
ncu -i gpurun_out/n_frame.ncu-rep --page raw --csv > gpurun_out/n_frame_raw.csv 2> /dev/null
rm -f gpurun_out/n_frame.ncu-rep        # 40 MB: gpurun_out travels back only below 64 MiB; the raw page is what the summary reads
# 3. build + refit at 10^7 triangles
timeout 600 python scripts/ncu_build.py > gpurun_out/n_build_plain.log 2>&1 && cat gpurun_out/n_build_plain.log && \
timeout 1500 $NCU --profile-from-start off --set full --import-source on -k regex:'k_fit|rs_scatter|rs_hist|k_emit_nodes|k_hierarchy|k_morton' -f -o gpurun_out/n_build python scripts/ncu_build.py > gpurun_out/n_build_ncu.log 2>&1
echo "build capture rc $?"
ncu -i gpurun_out/n_build.ncu-rep --page raw --csv > gpurun_out/n_build_raw.csv 2> /dev/null
rm -f gpurun_out/n_build.ncu-rep
timeout 600 python scripts/c2_parity_probe.py 2>&1 | tee gpurun_out/n_c2_parity.txt
ls -la gpurun_out | grep " n_"
