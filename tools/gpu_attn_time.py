"""Development aid: CUDA-event timing of the tensor-core attention forward at bench size (32 x 416^2, 6 heads,
unshifted and shifted) plus its error against the fp32 FMA-pipe kernel at batch 2; used to A/B library
variants (B200MAR_LIB=... python tools/gpu_attn_time.py)."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import transmar_gan_b200 as T
from oracle import ngswin_oracle as O
ops = T.ops
B = int(sys.argv[1]) if len(sys.argv) > 1 else 32
wname = sys.argv[2] if len(sys.argv) > 2 else "W1"
sd = {k: v.cuda() for k, v in O.synthetic_state_dict(wname, seed=1).items()}
P = [sd["encoder_layer1.blocks.1." + n].contiguous() for n in ops.NSTB_PARAM_NAMES]
ph = pw = 416
g = torch.Generator(device="cuda").manual_seed(5)
z = torch.randn(B, ph * pw, 64, device="cuda", generator=g)
c = torch.randn(B, (ph // 8) * (pw // 8), 64, device="cuda", generator=g) * 0.5
out = [os.path.basename(T.LIB_PATH)]
for shift in (0, 4):
    for _ in range(3):
        ops.attn_tc_fwd(z, c, *P[10:18], ph, pw, 6, shift)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    n = 20
    e0.record()
    for _ in range(n):
        ops.attn_tc_fwd(z, c, *P[10:18], ph, pw, 6, shift)
    e1.record()
    torch.cuda.synchronize()
    out.append(f"shift {shift}: {e0.elapsed_time(e1) / n:.4f} ms")
    a = ops.attn_tc_fwd(z[:2], c[:2], *P[10:18], ph, pw, 6, shift)
    r = ops.attn_fwd(z[:2].contiguous(), c[:2].contiguous(), *P[10:18], ph, pw, 6, shift)
    d = (a - r).abs()
    out.append(f"err max {d.max().item():.3e} mean {d.mean().item():.3e}")
print(" | ".join(out), "| tc_error", ops.tc_error())
