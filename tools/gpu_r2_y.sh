#!/bin/bash
mkdir -p gpurun_out
timeout 2400 python -m pytest tests -m gpu -x -q > gpurun_out/y_tests.log 2>&1; echo "all gpu tests rc=$?"; tail -3 gpurun_out/y_tests.log | cut -c1-200
python bench.py --no-train --no-cpu --no-highres --steps 10 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print('W1', round(d['value'],1), round(d['ms_per_step'],2))"
ncu --metrics gpu__time_duration.sum --clock-control none --profile-from-start off --csv --log-file gpurun_out/y_launches.csv python tools/gpu_fwd_once.py > /dev/null 2>&1; python tools/launch_summary.py gpurun_out/y_launches.csv | head -16
