#!/bin/bash
O=gpurun_out/k; mkdir -p $O
python tools/gpu_train_prof.py 16 > $O/train_prof.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none --profile-from-start off -c 4000 --csv --log-file $O/launches_train.csv \
    python tools/gpu_train_prof.py 16 > $O/ncu_train.log 2>&1
cat $O/train_prof.log
