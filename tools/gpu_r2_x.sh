#!/bin/bash
mkdir -p gpurun_out
timeout 2400 python -m pytest tests -m gpu -x -q > gpurun_out/x_tests.log 2>&1; echo "all gpu tests rc=$?"; tail -3 gpurun_out/x_tests.log | cut -c1-200
python - <<'PY'
import torch, transmar_gan_b200 as T
lib = T._cabi.lib()
B, H, W = 32, 416, 416
x = torch.rand(B, 1, H, W, device="cuda"); w = torch.randn(64, 1, 3, 3, device="cuda"); b = torch.randn(64, device="cuda")
c0 = torch.empty(B, H * W, 64, device="cuda")
st = torch.cuda.current_stream().cuda_stream
def t(f, n=10):
    for _ in range(3): f()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n): f()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n
print("shallow_tok", round(t(lambda: lib.b200mar_shallow_tok_fwd(x.data_ptr(), w.data_ptr(), b.data_ptr(), c0.data_ptr(), B, H, W, st)), 4), "ms")
PY
python bench.py --no-train --no-cpu --no-highres --steps 10 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print('W1', round(d['value'],1), round(d['ms_per_step'],2))"
