#!/bin/bash
# round 2, run B: the no-alloc test alone in a fresh process, then the whole GPU suite (it collects last), then bench
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_zz_no_alloc.py -m gpu -x -q > gpurun_out/b_noalloc_alone.log 2>&1
echo "alone rc $?" >> gpurun_out/b_noalloc_alone.log
tail -4 gpurun_out/b_noalloc_alone.log
timeout 1500 python -m pytest tests -m gpu -q > gpurun_out/b_pytest.log 2>&1
echo "pytest rc $?" >> gpurun_out/b_pytest.log
tail -30 gpurun_out/b_pytest.log
