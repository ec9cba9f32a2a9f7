"""Development aid: a few launches of the tensor-core backward entry points at the training size (for ncu)."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import transmar_gan_b200 as T
from oracle import ngswin_oracle as O
ops = T.ops
torch.manual_seed(0)
sd = {k: v.cuda() for k, v in O.synthetic_state_dict("W1", seed=1).items()}
B, ph, pw = int(sys.argv[1]) if len(sys.argv) > 1 else 16, 128, 128
for pre, nh, div in (("encoder_layer1.blocks.1.", 6, 1), ("encoder_layer2.blocks.1.", 4, 2)):
    P = [sd[pre + n].contiguous() for n in ops.NSTB_PARAM_NAMES]
    h, w = ph // div, pw // div
    z = torch.randn(B, h * w, 64, device="cuda")
    c = torch.randn(B, (h // 8) * (w // 8), 64, device="cuda") * 0.5
    da = torch.randn_like(z)
    dz = torch.empty_like(z)
    dctx = torch.zeros_like(c)
    G = [torch.zeros_like(p) for p in P]
    for it in range(3):
        ops.attn_bwd_tc(z, c, da, P[10:18], dz, dctx, G[10:18], h, w, nh, 4)
        ops.mlp_bwd_tc(z.view(-1, 64), da.view(-1, 64), P[18:24], dz.view(-1, 64), G[18:24])
torch.cuda.synchronize()
print("tc_error", ops.tc_error())
