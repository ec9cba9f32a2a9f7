#!/bin/bash
# full GPU suite as the round end runs it (-x), then the no-alloc test alone in a fresh process
mkdir -p gpurun_out
timeout 2400 python -m pytest tests -m gpu -x -q -s > gpurun_out/d_gpu_tests.log 2>&1; echo "gpu_tests rc=$?"
tail -4 gpurun_out/d_gpu_tests.log
grep -h "convergence\]\|\[auto/train\]\|W3 phantom\|W3 plan" gpurun_out/d_gpu_tests.log | tail -8
timeout 600 python -m pytest tests/test_zz_no_alloc.py -m gpu -x -q 2>&1 | tail -2
