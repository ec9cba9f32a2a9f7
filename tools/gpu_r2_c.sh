#!/bin/bash
mkdir -p gpurun_out
for shp in "1 64 64" "1 100 70" "2 96 160"; do timeout 600 python tools/gpu_repeat_check.py $shp 2>&1 | grep -v Warn; done | tee gpurun_out/c_repeat.log
