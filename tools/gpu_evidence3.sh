#!/bin/bash
# ncu --set full captures of the kernels added / rewritten at the end of round 1 (each program has already exited 0
# without ncu in this round's earlier runs; they are re-run plainly first anyway).
set -x
O=gpurun_out/ev3
mkdir -p $O
python tools/gpu_radon_time.py > $O/radon_time.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:radon_fwd -s 6 -c 1 -o $O/radon_fwd -f \
    python tools/gpu_radon_time.py > $O/ncu_radon_fwd.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:radon_bwd -s 6 -c 1 -o $O/radon_bwd -f \
    python tools/gpu_radon_time.py > $O/ncu_radon_bwd.log 2>&1
python tools/gpu_train_prof.py 16 > $O/train_prof.log 2>&1 && \
ncu --set full --clock-control none --import-source on --profile-from-start off -k regex:ngram_ctx_bwd_kernel -s 2 -c 1 -o $O/ngram_ctx_bwd -f \
    python tools/gpu_train_prof.py 16 > $O/ncu_ngram.log 2>&1
ncu --set full --clock-control none --import-source on --profile-from-start off -k regex:unigram_bwd -s 2 -c 1 -o $O/unigram_bwd -f \
    python tools/gpu_train_prof.py 16 > $O/ncu_unigram.log 2>&1
ls -la $O
