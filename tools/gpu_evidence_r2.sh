#!/bin/bash
# Round-2 evidence in one gpurun call: GPU tests, smoke, both bench arms, ncu launch lists (inference bench command
# and the training step) and ncu --set full captures of the two dominant kernels.  Every ncu command runs only
# after the same command has exited 0 without ncu.
set -x
O=gpurun_out/ev2
mkdir -p $O
python -m pytest tests -m gpu -x -q > $O/tests.log 2>&1; tail -2 $O/tests.log
python -c "import __graft_entry__ as g; g.smoke()" > $O/smoke.log 2>&1; grep "smoke\]" $O/smoke.log
python bench.py --impl reference --steps 2 --warmup 1 > $O/bench_reference.json 2> $O/bench_reference.err
python bench.py > $O/bench.json 2> $O/bench.err || exit 1
python bench.py --steps 2 --warmup 3 --no-train --no-cpu --no-highres > $O/bench_short.json 2>/dev/null && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 1800 --csv --log-file $O/launches_tc_b32.csv \
    python bench.py --steps 2 --warmup 3 --no-train --no-cpu --no-highres > $O/ncu_launches.log 2>&1
python tools/gpu_train_prof.py 16 > $O/train_prof.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none --profile-from-start off -c 4000 --csv --log-file $O/launches_train_b16x128.csv \
    python tools/gpu_train_prof.py 16 > $O/ncu_train.log 2>&1
python tools/gpu_attn_prof.py 32 > $O/attn_prof.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:attn_tc_fwd -s 2 -c 1 -o $O/attn6 -f \
    python tools/gpu_attn_prof.py 32 > $O/ncu_attn.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:mlp_tma_fwd -s 2 -c 1 -o $O/mlp_tma -f \
    python tools/gpu_attn_prof.py 32 > $O/ncu_mlp.log 2>&1
ls -la $O
