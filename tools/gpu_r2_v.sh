#!/bin/bash
mkdir -p gpurun_out
timeout 1800 python -m pytest tests/test_gpu_parity.py tests/test_gpu_tf32.py tests/test_gpu_tf32c.py tests/test_gpu_auto_precision.py -m gpu -x -q 2>&1 | tail -3 | cut -c1-200
python tools/gpu_s16_time.py 2>&1 | grep attn
python bench.py --no-train --no-cpu --no-highres --steps 10 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print('W1', round(d['value'],1), round(d['ms_per_step'],2), 'attn live', d['roofline']['avg_launch_ms'], d['roofline']['frac'])"
