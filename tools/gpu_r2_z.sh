#!/bin/bash
mkdir -p gpurun_out
echo "--- TMA"; python tools/gpu_mlp_bwd_ab.py 2>&1 | tail -6
echo "--- plain"; B200MAR_MLP_BWD_PLAIN=1 python tools/gpu_mlp_bwd_ab.py 2>&1 | tail -6
timeout 2400 python -m pytest tests/test_gpu_bwd_tc.py tests/test_gpu_train.py tests/test_gpu_tf32.py tests/test_gpu_guard_bands.py -m gpu -x -q 2>&1 | tail -3 | cut -c1-200
