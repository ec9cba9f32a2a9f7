#!/bin/bash
# End-of-round evidence refresh: GPU tests, the default bench line, reference arm, smoke, and the ncu launch list of
# the full GAN step (fused loss kernel + Radon kernels); the attention / MLP captures of gpu_evidence.sh stand (those
# kernels did not change).  Every ncu command runs only after the same command has exited 0 without ncu.
set -x
O=gpurun_out/ev
mkdir -p $O
python -m pytest tests -m gpu -x -q > $O/tests.log 2>&1; tail -2 $O/tests.log
python bench.py > $O/bench.json 2> $O/bench.err || exit 1
python bench.py --impl reference --steps 2 --warmup 1 > $O/bench_reference.json 2> $O/bench_reference.err
python -c "import __graft_entry__ as g; g.smoke()" > $O/smoke.log 2>&1; tail -3 $O/smoke.log
python tools/gpu_gan_prof.py > $O/gan_prof.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none --profile-from-start off -c 6000 --csv --log-file $O/launches_gan_b16x128.csv \
    python tools/gpu_gan_prof.py > $O/ncu_gan.log 2>&1
cat $O/gan_prof.log
ls -la $O
