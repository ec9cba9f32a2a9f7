"""Development aid: bottleneck tail tc vs fp32 kernel + timing."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import transmar_gan_b200 as T
from oracle import ngswin_oracle as O
ops = T.ops
def run(prec, args):
    ops.set_precision(prec)
    return ops.bottleneck_fwd(*args)
for kind in ("W1", "W2"):
    sd = {k: v.cuda() for k, v in O.synthetic_state_dict(kind, seed=1).items()}
    W = [sd["bottleneck." + n].contiguous() for n in ("depthwise.weight", "depthwise.bias", "pointwise.weight", "pointwise.bias", "norm.weight", "norm.bias")]
    for (B, H, Wd) in ((1, 16, 16), (2, 24, 40), (1, 64, 64), (2, 104, 72)):
        torch.manual_seed(H)
        mk = lambda h, w: torch.randn(B, h * w, 64, device="cuda")
        args = [mk(H, Wd), mk(H // 2, Wd // 2), mk(H // 4, Wd // 4), mk(H, Wd), mk(H // 2, Wd // 2), mk(H // 4, Wd // 4), *W, H, Wd]
        ref = run("fp32", args); got = run("tf32", args)
        torch.cuda.synchronize()
        d = (ref - got).abs()
        print(kind, B, H, Wd, "max", d.max().item(), "mean", d.mean().item(), "ref absmax", ref.abs().max().item(), "tc_error", ops.tc_error(), flush=True)
def timeit(f, n=5):
    f(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n): f()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n
B, H, Wd = 32, 416, 416
mk = lambda h, w: torch.randn(B, h * w, 64, device="cuda")
args = [mk(H, Wd), mk(H // 2, Wd // 2), mk(H // 4, Wd // 4), mk(H, Wd), mk(H // 2, Wd // 2), mk(H // 4, Wd // 4), *W, H, Wd]
print("fp32 cat+tail ms", timeit(lambda: run("fp32", args)))
print("tc   cat+tail ms", timeit(lambda: run("tf32", args)), "tc_error", ops.tc_error())
