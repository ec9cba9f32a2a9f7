"""Development aid: cascade_proj tc vs fp32 kernel + timing; also times shallow / final kernels."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import transmar_gan_b200 as T
from oracle import ngswin_oracle as O
ops = T.ops
def run(prec, *a):
    ops.set_precision(prec)
    return ops.cascade_proj_fwd(*a)
for kind in ("W1", "W2"):
    sd = {k: v.cuda() for k, v in O.synthetic_state_dict(kind, seed=1).items()}
    for lay, two in (("encoder_layer2", False), ("encoder_layer3", True)):
        w, bias = sd[lay + ".across_cascade_proj.weight"].contiguous(), sd[lay + ".across_cascade_proj.bias"].contiguous()
        for (B, h, wd) in ((1, 8, 16), (2, 12, 20), (1, 32, 32), (2, 52, 36)):
            torch.manual_seed(h)
            A = torch.randn(B, 4 * h * wd, 64, device="cuda")
            Bs = torch.randn(B, 4 * h * wd, 64, device="cuda") if two else None
            D = torch.randn(B, h * wd, 64, device="cuda")
            ro, rp = run("fp32", A, Bs, D, w, bias, h, wd)
            go, gp = run("tf32", A, Bs, D, w, bias, h, wd)
            torch.cuda.synchronize()
            d = (ro - go).abs()
            print(kind, lay, B, h, wd, "max", d.max().item(), "mean", d.mean().item(), "pool equal", torch.equal(rp, gp), "tc_error", ops.tc_error(), flush=True)
def timeit(f, n=5):
    f(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n): f()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n
B, h, wd = 32, 208, 208
sd = {k: v.cuda() for k, v in O.synthetic_state_dict("W1", seed=1).items()}
w, bias = sd["encoder_layer2.across_cascade_proj.weight"].contiguous(), sd["encoder_layer2.across_cascade_proj.bias"].contiguous()
A = torch.randn(B, 4 * h * wd, 64, device="cuda"); D = torch.randn(B, h * wd, 64, device="cuda")
print("cascade fp32 ms", timeit(lambda: run("fp32", A, None, D, w, bias, h, wd)))
print("cascade tc   ms", timeit(lambda: run("tf32", A, None, D, w, bias, h, wd)), "tc_error", ops.tc_error())
x = torch.randn(B, 1, 416, 416, device="cuda")
se = (sd["shallow_extract.conv1.weight"].contiguous(), sd["shallow_extract.conv1.bias"].contiguous())
print("shallow ms", timeit(lambda: ops.shallow_tok_fwd(x, *se)))
tt = [sd[n].contiguous() for n in ("norm.weight", "norm.bias", "to_target.before_shuffle.weight", "to_target.before_shuffle.bias", "to_target.to_origin.weight", "to_target.to_origin.bias")]
print("final recon ms", timeit(lambda: ops.final_recon_fwd(A, A, *tt, 416, 416, 416, 416)))
