#!/bin/bash
# training-path confirmation: backward tests + eager L1 step timing + graph step timing
mkdir -p gpurun_out
python -m pytest tests/test_gpu_bwd_tc.py tests/test_gpu_train.py tests/test_gpu_tf32.py -m gpu -x -q 2>&1 | tail -3 | cut -c1-300
python tools/gpu_train_prof.py 16 2>&1 | tail -1
python tools/gpu_train_graph.py 2>&1 | tail -4
