#!/bin/bash
mkdir -p gpurun_out
timeout 2400 python -m pytest tests -m gpu -x -q > gpurun_out/u_tests.log 2>&1; echo "all gpu tests rc=$?"; tail -3 gpurun_out/u_tests.log | cut -c1-200
python bench.py --no-cpu --steps 10 2>/dev/null > gpurun_out/u_bench.json; python -c "
import json; d=json.loads(open('gpurun_out/u_bench.json').read())
for k in ('trained_like','stress_set'):
    t=d[k]; print(k, round(t['slices_per_s'],1), round(t['ms_per_step'],2), t['units_on_tensor_cores'], t['plan'])
print('W1', round(d['value'],1), 'e2e', round(d['e2e']['value'],1), 'bf16', round(d['bf16_stream']['slices_per_s'],1), 'highres', d['highres'].get('slices_per_s'), 'train', d['train']['ms_per_step'], 'gan', d['train_gan']['ms_per_step'], 'frac', d['roofline']['frac'])"
