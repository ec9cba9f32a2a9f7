#!/bin/bash
mkdir -p gpurun_out
timeout 900 python tools/gpu_plan_report.py > gpurun_out/b_plan.log 2>&1; echo "plan rc=$?"
grep -v Warn gpurun_out/b_plan.log | tail -30
timeout 2400 python -m pytest tests -m gpu -q -s > gpurun_out/b_gpu_tests.log 2>&1; echo "gpu_tests rc=$?"
tail -15 gpurun_out/b_gpu_tests.log
grep -h "\[auto\]\|\[convergence\]\|\[auto/train\]" gpurun_out/b_gpu_tests.log | tail -12
for lib in "" /root/repo/transmar_gan_b200/libb200mar_w3.so; do
  echo "== bench lib=${lib:-product}"
  B200MAR_LIB=$lib timeout 600 python bench.py --no-train --no-cpu --no-highres --steps 10 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read()); r=d['roofline']; print('slices/s', round(d['value'],1), 'ms/step', round(d['ms_per_step'],2), 'attn6 ms', round(r['avg_launch_ms'],4), 'mlp ms', round(r['mlp']['avg_launch_ms'],4), 'e2e', round(d['e2e']['value'],1), d['clocks'])"
done
