#!/bin/bash
mkdir -p gpurun_out
timeout 1200 python -m pytest tests/test_gpu_tf32c.py -m gpu -x -q -s 2>&1 | grep "tf32c\] mlp\|tf32c\] W\|passed\|failed\|Error\|error\|assert" | tail -40 | cut -c1-250
python - <<'PY'
import torch, transmar_gan_b200 as T
from oracle import ngswin_oracle as O
ops = T.ops
sd = {k: v.cuda() for k, v in O.synthetic_state_dict("W1", seed=1).items()}
P = [sd["encoder_layer1.blocks.1." + n].contiguous() for n in ops.NSTB_PARAM_NAMES]
n = 32 * 416 * 416
a = torch.randn(1, n, 64, device="cuda"); z = torch.randn(1, n, 64, device="cuda")
for comp in (False, True):
    for _ in range(2): ops.mlp_tma_fwd_(a, z, *P[18:24], False, comp)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(5): ops.mlp_tma_fwd_(a, z, *P[18:24], False, comp)
    e1.record(); torch.cuda.synchronize()
    print("mlp_tma comp" if comp else "mlp_tma plain", round(e0.elapsed_time(e1) / 5, 3), "ms")
print("tc_error", ops.tc_error())
PY
