#!/bin/bash
mkdir -p gpurun_out
timeout 1800 python -m pytest tests/test_gpu_auto_precision.py tests/test_gpu_parity.py tests/test_gpu_train.py tests/test_gpu_bwd_tc.py tests/test_gpu_tf32.py -m gpu -x -q -s > gpurun_out/l_tests.log 2>&1; echo "tests rc=$?"; tail -3 gpurun_out/l_tests.log | cut -c1-200
grep -h "W3 plan\|W3 phantom\|\[auto\] W2 (1, 64, 64)" gpurun_out/l_tests.log | head -4
python tools/gpu_plan_report.py 2>&1 | grep -v Warn | grep "plan:\|all-tc" 
python bench.py --no-train --no-cpu --steps 5 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read()); t=d['trained_like']; print('W1', round(d['value'],1), '| W3', round(t['slices_per_s'],1), t['plan'], t['units_on_tensor_cores'], t['probe_err_plan'])"
