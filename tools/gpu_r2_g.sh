#!/bin/bash
mkdir -p gpurun_out
timeout 2400 python -m pytest tests -m gpu -x -q -s > gpurun_out/g_gpu_tests.log 2>&1; echo "gpu_tests rc=$?"
tail -3 gpurun_out/g_gpu_tests.log | cut -c1-200
grep -h "^\.*\[auto\] W[123] (1, 100, 70)\|W3 phantom\|W3 plan\|convergence\] loss" gpurun_out/g_gpu_tests.log | tail -6
timeout 900 python bench.py --no-train --no-cpu --no-highres --steps 10 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read()); r=d['roofline']; print('slices/s', round(d['value'],1), 'ms/step', round(d['ms_per_step'],2), 'attn6 ms', round(r['avg_launch_ms'],4), 'mlp ms', round(r['mlp']['avg_launch_ms'],4), 'e2e', round(d['e2e']['value'],1), 'bf16', round(d['bf16_stream']['slices_per_s'],1), d['clocks']['reasons'])"
