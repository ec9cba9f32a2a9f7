#!/bin/bash
# quick confirmation of a change: attention timing + the whole GPU test suite
mkdir -p gpurun_out
python tools/gpu_attn_time.py 32 2>&1 | tail -1 | tee gpurun_out/check_attn.log
python -m pytest tests -m gpu -x -q 2>&1 | tail -4 | cut -c1-300 | tee gpurun_out/check_tests.log
