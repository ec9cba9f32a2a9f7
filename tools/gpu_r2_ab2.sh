#!/bin/bash
# tools/gpu_r2_ab2.sh <script.py> <tag> <tag> ... : run a timing script against library variants ("base" = product)
mkdir -p gpurun_out
s=$1; shift
for t in "$@"; do
  if [ "$t" = base ]; then lib=libb200mar.so; else lib=libb200mar_$t.so; fi
  B200MAR_LIB=$(pwd)/transmar_gan_b200/$lib python $s 2>&1 | tail -1
done | tee gpurun_out/ab2.log
