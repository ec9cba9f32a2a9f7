"""Scratch GPU check of the tensor-core kernels against the fp32 SIMT kernels (development aid)."""
import sys, os, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import transmar_gan_b200 as T
from oracle import ngswin_oracle as O

ops = T.ops
torch.manual_seed(0)
for kind in ("W1", "W2"):
    sd = {k: v.cuda() for k, v in O.synthetic_state_dict(kind, seed=1).items()}
    pre = "encoder_layer1.blocks.0."
    P = [sd[pre + n].contiguous() for n in ops.NSTB_PARAM_NAMES][18:24]
    for ntok in (64, 128, 192, 64 * 40, 64 * 2704):
        a = torch.randn(ntok, 64, device="cuda")
        z = torch.randn(ntok, 64, device="cuda")
        o1, z1 = ops.mlp_fwd(a, z, *P, True, True)
        o2, z2 = ops.mlp_tc_fwd(a, z, *P, True, True)
        torch.cuda.synchronize()
        print(kind, ntok, "mlp tc vs simt: out", (o1 - o2).abs().max().item(), "zsum",
              (z1 - z2).abs().max().item(), "scale", o1.abs().max().item(), "tc_error", ops.tc_error())
a = torch.randn(32 * 416 * 416, 64, device="cuda")
z = torch.randn_like(a)
for name, fn in (("simt", ops.mlp_fwd), ("tc", ops.mlp_tc_fwd)):
    for _ in range(2):
        fn(a, z, *P, False, True)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(5):
        fn(a, z, *P, False, True)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 5
    gb = a.numel() * 4 * 3 / 1e9
    print(f"mlp {name}: {ms:.3f} ms per launch (B=32 416^2), {gb / ms * 1e3:.0f} GB/s algorithmic (3 passes)")
