#!/bin/bash
mkdir -p gpurun_out
timeout 1800 python -m pytest tests/test_gpu_parity.py tests/test_gpu_auto_precision.py tests/test_gpu_guard_bands.py -m gpu -x -q > gpurun_out/m_tests.log 2>&1; echo "tests rc=$?"; tail -3 gpurun_out/m_tests.log | cut -c1-200
python - <<'PY'
import torch, transmar_gan_b200 as T
from oracle import ngswin_oracle as O
ops, lib = T.ops, T._cabi.lib()
B, ph, pw, nh = 32, 416, 416, 6
sd = {k: v.cuda() for k, v in O.synthetic_state_dict("W1", seed=1).items()}
P = [sd["encoder_layer1.blocks.1." + n].contiguous() for n in ops.NSTB_PARAM_NAMES]
p = [t.data_ptr() for t in P]
z = torch.randn(B, ph * pw, 64, device="cuda"); a = torch.empty_like(z)
ctx = torch.randn(B, (ph // 8) * (pw // 8), 64, device="cuda") * 0.5
st = torch.cuda.current_stream().cuda_stream
for shift in (0, 4):
    for _ in range(2): lib.b200mar_attn_fwd(z.data_ptr(), ctx.data_ptr(), *p[10:18], a.data_ptr(), B, ph, pw, nh, shift, st)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(3): lib.b200mar_attn_fwd(z.data_ptr(), ctx.data_ptr(), *p[10:18], a.data_ptr(), B, ph, pw, nh, shift, st)
    e1.record(); torch.cuda.synchronize()
    print("attn_fwd fp32 shift", shift, round(e0.elapsed_time(e1) / 3, 3), "ms")
PY
python bench.py --no-train --no-cpu --steps 5 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read()); t=d['trained_like']; print('W1', round(d['value'],1), '| W3', round(t['slices_per_s'],1), round(t['ms_per_step'],2), t['units_on_tensor_cores'])"
