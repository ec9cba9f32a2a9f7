"""Development aid: patch_merge_tc_fwd vs patch_merge_fwd + timing."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import transmar_gan_b200 as T
from oracle import ngswin_oracle as O
ops = T.ops
for kind in ("W1", "W2"):
    sd = {k: v.cuda() for k, v in O.synthetic_state_dict(kind, seed=1).items()}
    lw, lb, rw = (sd["encoder_layer1.downsample." + n].contiguous() for n in ("norm.weight", "norm.bias", "reduction.weight"))
    for (B, ph, pw) in ((1, 16, 32), (2, 24, 40), (1, 64, 64), (3, 104, 104)):
        torch.manual_seed(ph)
        x = torch.randn(B, ph * pw, 64, device="cuda") * 2 + 0.7
        ref = ops.patch_merge_fwd(x, lw, lb, rw, ph, pw)
        got = ops.patch_merge_tc_fwd(x, lw, lb, rw, ph, pw)
        torch.cuda.synchronize()
        d = (ref - got).abs()
        print(kind, B, ph, pw, "max", d.max().item(), "mean", d.mean().item(), "ref absmax", ref.abs().max().item(), "tc_error", ops.tc_error(), flush=True)
def timeit(f, n=5):
    f(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n): f()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n
x = torch.randn(32, 416 * 416, 64, device="cuda")
print("simt ms", timeit(lambda: ops.patch_merge_fwd(x, lw, lb, rw, 416, 416)))
print("tc   ms", timeit(lambda: ops.patch_merge_tc_fwd(x, lw, lb, rw, 416, 416)), "tc_error", ops.tc_error())
