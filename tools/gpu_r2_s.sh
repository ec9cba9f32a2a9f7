#!/bin/bash
mkdir -p gpurun_out
timeout 2400 python -m pytest tests -m gpu -x -q > gpurun_out/s_tests.log 2>&1; echo "all gpu tests rc=$?"; tail -3 gpurun_out/s_tests.log | cut -c1-200
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | grep "smoke\]" | cut -c1-250
( time python bench.py --no-cpu --steps 5 2>gpurun_out/s_bench.err > gpurun_out/s_bench.json ) 2>&1 | grep real; python -c "
import json; d=json.loads(open('gpurun_out/s_bench.json').read())
for k in ('trained_like','stress_set'):
    t=d[k]; print(k, round(t['slices_per_s'],1), round(t['ms_per_step'],2), t['units_on_tensor_cores'], t['plan'])
print('W1', round(d['value'],1), 'train', d['train']['ms_per_step'], 'gan', d['train_gan']['ms_per_step'])"
