"""Development aid: per-phase cycle profile of attn_bwd_core_kernel (library built with
B200MAR_NVCC_EXTRA=-DB200MAR_ABWD_PROF)."""
import sys, os, ctypes
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import transmar_gan_b200 as T
from transmar_gan_b200 import _cabi
from oracle import ngswin_oracle as O
ops = T.ops
L = _cabi.lib()
sd = {k: v.cuda() for k, v in O.synthetic_state_dict("W1", seed=1).items()}
P = [sd["encoder_layer1.blocks.1." + n].contiguous() for n in ops.NSTB_PARAM_NAMES][10:18]
B, ph, pw = 16, 128, 128
z = torch.randn(B, ph * pw, 64, device="cuda"); c = torch.randn(B, 256, 64, device="cuda") * 0.5
da = torch.randn_like(z); dz = torch.empty_like(z); dctx = torch.zeros_like(c)
G = [torch.zeros_like(p) for p in P]
_, o16 = ops.attn_tc_fwd_save(z, c, *P, ph, pw, 6, 4)
buf = (ctypes.c_ulonglong * 32)()
ops.attn_bwd_tc(z, c, da, P, dz, dctx, G, ph, pw, 6, 4, o16); torch.cuda.synchronize()
L.b200mar_abwd_prof(buf, 1)
ops.attn_bwd_tc(z, c, da, P, dz, dctx, G, ph, pw, 6, 4, o16); torch.cuda.synchronize()
L.b200mar_abwd_prof(buf, 1)
tot = sum(buf)
names = ["loop", "X tile", "O/Y/LN/dO", "qkv+norm+tiles", "dO_h ops", "sync", "S+dP mma", "softmax", "dP/rowdot/PT",
         "dS+scatter", "dS operands", "sync", "dV/dq/dk mma", "normbwd+export", "sync", "bias-table grad"]
for i, n in enumerate(names):
    print(f"{i:2d} {n:18s} {buf[i]:10d} cycles {100.0 * buf[i] / tot:5.1f}%")
print("total cycles", tot, "tc_error", ops.tc_error())
