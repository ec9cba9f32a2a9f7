#!/bin/bash
mkdir -p gpurun_out
timeout 2400 python -m pytest tests -m gpu -x -q > gpurun_out/j_tests.log 2>&1; echo "tests rc=$?"; tail -3 gpurun_out/j_tests.log | cut -c1-200
timeout 900 python bench.py --no-cpu --no-highres --steps 5 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read());
for k in ('train','train_gan'):
    t=d[k]; print(k, round(t['ms_per_step'],2), 'ms graph', round(t['graph_ms_per_step'],2), 'eager', round(t['eager_ms_per_step'],2), 'launches', t['gpu_launches_per_step'], 'loss', t.get('final_loss', t.get('final_loss_G')))
print('gan reuse', round(d['train_gan']['reuse_fake_ms_per_step'],2), 'phys', round(d['train_gan']['with_physics_ms_per_step'],2), '| inference', round(d['value'],1))"
