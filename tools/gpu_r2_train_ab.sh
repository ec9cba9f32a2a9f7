#!/bin/bash
# A/B of library variants on the L1 training step (graph replay) : tools/gpu_r2_train_ab.sh <tag> <tag> ...
mkdir -p gpurun_out
for t in "$@"; do
  if [ "$t" = base ]; then lib=libb200mar.so; else lib=libb200mar_$t.so; fi
  echo "== $lib"
  B200MAR_LIB=$(pwd)/transmar_gan_b200/$lib python tools/gpu_train_graph.py 2>&1 | grep -E "graph|eager" | head -2
done | tee gpurun_out/train_ab.log
