#!/bin/bash
mkdir -p gpurun_out
python tools/gpu_plan_report.py > gpurun_out/n_plan.log 2>&1; echo "plan rc=$?"; grep -v "unit 0:" gpurun_out/n_plan.log | cut -c1-400
timeout 1800 python -m pytest tests/test_gpu_auto_precision.py tests/test_gpu_parity.py -m gpu -x -q > gpurun_out/n_tests.log 2>&1; echo "tests rc=$?"; tail -3 gpurun_out/n_tests.log | cut -c1-200
python - <<'PY'
import time, torch, transmar_gan_b200 as T
from oracle import ngswin_oracle as O
m = T.NGswin(); m.load_state_dict(O.synthetic_state_dict("W2", seed=1)); m = m.cuda().eval()
m.calibrate_precision(); torch.cuda.synchronize()
t = time.time(); r = m.calibrate_precision(); torch.cuda.synchronize(); print("W2 calibrate", round(time.time() - t, 2), "s")
m = T.NGswin(); m.load_state_dict(O.synthetic_state_dict("W1", seed=1)); m = m.cuda().eval()
t = time.time(); r = m.calibrate_precision(); torch.cuda.synchronize(); print("W1 calibrate", round(time.time() - t, 2), "s")
PY
python bench.py --no-train --no-cpu --steps 5 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read()); t=d['trained_like']; print('W1', round(d['value'],1), '| W3', round(t['slices_per_s'],1), round(t['ms_per_step'],2), t['units_on_tensor_cores'])"
