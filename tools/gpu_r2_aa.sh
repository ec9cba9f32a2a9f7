#!/bin/bash
mkdir -p gpurun_out
timeout 2400 python -m pytest tests -m gpu -x -q > gpurun_out/aa_tests.log 2>&1; echo "all gpu tests rc=$?"; tail -3 gpurun_out/aa_tests.log | cut -c1-200
python bench.py --no-cpu --no-highres --steps 10 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print('W1', round(d['value'],1), 'train', round(d['train']['ms_per_step'],3), 'eager', round(d['train']['eager_ms_per_step'],2), 'gan', round(d['train_gan']['ms_per_step'],3))"
