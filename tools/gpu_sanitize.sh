#!/bin/bash
# usage: gpu_sanitize.sh memcheck|racecheck   - ONE tool per gpurun call (B200_PROFILING.md), smallest cases
tool=$1
mkdir -p gpurun_out
K='(attn_and_mlp_tc_forward and W1 and geom3) or (bf16_stream_kernels and W1 and geom3) or no_alloc_entry_point_matches or mlp_tma_inplace or mlp_bwd_tc_vs_fp32 or wgrad_tc_vs_matmul or attn_bwd_tc_saved_o or cascade_proj_tc or patch_merge_tc or bottleneck_tc_matches'
python -m pytest tests/test_gpu_tf32.py tests/test_gpu_bf16_stream.py tests/test_zz_no_alloc.py tests/test_gpu_bwd_tc.py -m gpu -q -k "$K" --collect-only 2>/dev/null | tail -3
timeout 1500 compute-sanitizer --tool $tool --print-limit 20 python -m pytest tests/test_gpu_tf32.py tests/test_gpu_bf16_stream.py tests/test_zz_no_alloc.py tests/test_gpu_bwd_tc.py -m gpu -q -x -k "$K" > gpurun_out/sanitize_$tool.log 2>&1
echo "sanitizer rc=$?"
grep -E "ERROR SUMMARY|RACECHECK SUMMARY|passed|failed|Error|hazard" gpurun_out/sanitize_$tool.log | sort | uniq -c | head -20
