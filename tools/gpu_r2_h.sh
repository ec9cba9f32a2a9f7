#!/bin/bash
for i in 1 2 3 4 5; do
  echo "== run $i"
  timeout 600 python -m pytest tests/test_gpu_convergence.py -m gpu -q -s 2>&1 | grep "convergence\] first\|passed\|failed"
done 2>&1 | tee gpurun_out/h_conv.log
