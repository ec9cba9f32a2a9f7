#!/bin/bash
mkdir -p gpurun_out
timeout 1200 python -m pytest tests/test_gpu_tf32c.py -m gpu -x -q -s 2>&1 | grep "tf32c\]\|passed\|failed\|Error\|error" | tail -30 | cut -c1-250
timeout 2400 python -m pytest tests -m gpu -x -q > gpurun_out/q_tests.log 2>&1; echo "all gpu tests rc=$?"; tail -3 gpurun_out/q_tests.log | cut -c1-200
python bench.py --no-cpu --steps 5 2>/dev/null > gpurun_out/q_bench.json; python -c "
import json; d=json.loads(open('gpurun_out/q_bench.json').read()); t=d['trained_like']; print('W1', round(d['value'],1), '| W3', round(t['slices_per_s'],1), round(t['ms_per_step'],2), t['units_on_tensor_cores'], t['plan']); print('train', d['train']['ms_per_step'], 'gan', d['train_gan']['ms_per_step'])"
