#!/bin/bash
mkdir -p gpurun_out
timeout 1200 python -m pytest tests/test_gpu_tf32c.py -m gpu -x -q -s 2>&1 | grep "tf32c\]\|passed\|failed\|Error\|error" | tail -25 | cut -c1-330
python tools/gpu_tf32c_probe.py 2>&1 | tail -20
python - <<'PY'
import torch, transmar_gan_b200 as T
from oracle import ngswin_oracle as O
ops = T.ops
B, ph, pw, nh = 32, 416, 416, 6
sd = {k: v.cuda() for k, v in O.synthetic_state_dict("W1", seed=1).items()}
P = [sd["encoder_layer1.blocks.1." + n].contiguous() for n in ops.NSTB_PARAM_NAMES]
z = torch.randn(B, ph * pw, 64, device="cuda")
ctx = torch.randn(B, (ph // 8) * (pw // 8), 64, device="cuda") * 0.5
for comp in (False, True):
    for shift in (0, 4):
        for _ in range(2): ops.attn_tc_fwd(z, ctx, *P[10:18], ph, pw, nh, shift, True, comp)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(5): ops.attn_tc_fwd(z, ctx, *P[10:18], ph, pw, nh, shift, True, comp)
        e1.record(); torch.cuda.synchronize()
        print("attn_tc comp" if comp else "attn_tc plain", "shift", shift, round(e0.elapsed_time(e1) / 5, 3), "ms")
print("tc_error", ops.tc_error())
PY
