"""Development aid: device time of the Radon forward / backprojection kernels at the training shape
(batch 16 x 128x128, 180 angles, 128 detectors), CUDA events around 50 launches each."""
import math, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from transmar_gan_b200 import radon

B, S, A = 16, 128, 180
proj = radon.Radon(S, torch.linspace(0, math.pi, A))
x = torch.rand(B, S, S, device="cuda")
y = torch.rand(B, A, S, device="cuda")
for name, fn, arg in (("forward", proj.forward, x), ("backprojection", proj.backprojection, y)):
    for _ in range(5):
        fn(arg)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(50):
        fn(arg)
    e1.record()
    torch.cuda.synchronize()
    print(f"radon {name}: {e0.elapsed_time(e1) / 50 * 1e3:.1f} us per launch (batch {B} x {S}x{S}, {A} angles)")
