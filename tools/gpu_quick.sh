#!/bin/bash
# Development aid: attention/NSTB parity tests + the inference bench line (attention and MLP launch averages).
python -m pytest tests/test_gpu_tf32.py tests/test_gpu_bwd_tc.py tests/test_gpu_parity.py -m gpu -x -q 2>&1 | tail -4
python bench.py --no-train --no-cpu --no-highres 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read()); r=d['roofline']; print('slices/s', round(d['value'],1), 'ms/step', round(d['ms_per_step'],2), 'attn6 ms', round(r['avg_launch_ms'],4), 'mlp ms', round(r['mlp']['avg_launch_ms'],4), 'e2e', round(d['e2e']['value'],1))"
