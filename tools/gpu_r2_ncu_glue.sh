#!/bin/bash
# ncu --set full captures of the glue backward kernels of one training step (after the same command ran clean)
mkdir -p gpurun_out/glue
python tools/gpu_train_prof.py 16 > gpurun_out/glue/prof.log 2>&1 || exit 1
for k in bottleneck_tail_bwd bottleneck_dw_bwd final_bwd patch_merge_bwd cascade_proj_bwd; do
  ncu --set full --clock-control none --import-source on --profile-from-start off -k regex:$k -c 1 -o gpurun_out/glue/$k -f \
      python tools/gpu_train_prof.py 16 > gpurun_out/glue/ncu_$k.log 2>&1
done
ls -la gpurun_out/glue
