#!/bin/bash
mkdir -p gpurun_out
timeout 900 python scripts/c4_walk.py > gpurun_out/g_c4_walk.txt 2>&1
echo "rc $?"
cat gpurun_out/g_c4_walk.txt | cut -c1-330
