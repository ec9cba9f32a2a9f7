#!/bin/bash
mkdir -p gpurun_out
timeout 1800 python -m pytest tests/test_gpu_parity.py tests/test_gpu_tf32.py tests/test_gpu_tf32c.py tests/test_gpu_bf16_stream.py tests/test_gpu_guard_bands.py -m gpu -x -q 2>&1 | tail -3 | cut -c1-200
python tools/gpu_s16_time.py 2>&1 | tail -12
python bench.py --no-train --no-cpu --steps 10 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read()); t=d['trained_like']; print('W1', round(d['value'],1), round(d['ms_per_step'],2), '| W3', round(t['slices_per_s'],1), '| bf16', round(d['bf16_stream']['slices_per_s'],1), '| attn', d['roofline'].get('kernel_ms'), d['roofline']['frac'])"
