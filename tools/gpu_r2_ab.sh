#!/bin/bash
# A/B of library variants: tools/gpu_r2_ab.sh <log name> <tag> <tag> ...   ("base" = the product library)
mkdir -p gpurun_out
log=$1; shift
for t in "$@"; do
  if [ "$t" = base ]; then lib=libb200mar.so; else lib=libb200mar_$t.so; fi
  B200MAR_LIB=$(pwd)/transmar_gan_b200/$lib python tools/gpu_attn_time.py 32 2>&1 | tail -1
done | tee gpurun_out/$log
