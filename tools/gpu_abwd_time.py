"""Development aid: CUDA-event timing of the training-size attention forward (save) and backward entry points
(16 x 128^2, 6 heads, shifted), for A/B runs of library variants (B200MAR_LIB=...)."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import transmar_gan_b200 as T
from oracle import ngswin_oracle as O
ops = T.ops
sd = {k: v.cuda() for k, v in O.synthetic_state_dict("W1", seed=1).items()}
P = [sd["encoder_layer1.blocks.1." + n].contiguous() for n in ops.NSTB_PARAM_NAMES][10:18]
B, ph, pw = 16, 128, 128
z = torch.randn(B, ph * pw, 64, device="cuda"); c = torch.randn(B, 256, 64, device="cuda") * 0.5
da = torch.randn_like(z); dz = torch.empty_like(z); dctx = torch.zeros_like(c)
G = [torch.zeros_like(p) for p in P]
_, o16 = ops.attn_tc_fwd_save(z, c, *P, ph, pw, 6, 4)


def t(f, n=50):
    for _ in range(5):
        f()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n):
        f()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n * 1e3


fw = t(lambda: ops.attn_tc_fwd_save(z, c, *P, ph, pw, 6, 4))
bw = t(lambda: ops.attn_bwd_tc(z, c, da, P, dz, dctx, G, ph, pw, 6, 4, o16))
print(f"{os.path.basename(T.LIB_PATH)} | fwd_save {fw:.1f} us | bwd (core + dx + 2 wgrad) {bw:.1f} us | tc_error {ops.tc_error()}")
