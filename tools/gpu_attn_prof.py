"""Development aid: a few launches of the tensor-core attention / MLP kernels at bench size, for ncu."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import transmar_gan_b200 as T
from oracle import ngswin_oracle as O
ops = T.ops
B = int(sys.argv[1]) if len(sys.argv) > 1 else 8
sd = {k: v.cuda() for k, v in O.synthetic_state_dict("W1", seed=1).items()}
P = [sd["encoder_layer1.blocks.1." + n].contiguous() for n in ops.NSTB_PARAM_NAMES]
ph = pw = 416
z = torch.randn(B, ph * pw, 64, device="cuda")
c = torch.randn(B, (ph // 8) * (pw // 8), 64, device="cuda") * 0.5
for shift in (0, 4, 0, 4):
    a = ops.attn_tc_fwd(z, c, *P[10:18], ph, pw, 6, shift)
    o = ops.mlp_tma_fwd_(a, z, *P[18:24], False)
torch.cuda.synchronize()
print("ok", ops.tc_error())
