"""Development aid: per-phase cycle profile of attn_tc (library built with -DB200MAR_ATTN_PROF)."""
import sys, os, ctypes
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import transmar_gan_b200 as T
from oracle import ngswin_oracle as O
ops = T.ops
L = ctypes.CDLL(T.LIB_PATH)
L.b200mar_attn_prof.argtypes = [ctypes.c_void_p, ctypes.c_int]
sd = {k: v.cuda() for k, v in O.synthetic_state_dict("W1", seed=1).items()}
P = [sd["encoder_layer1.blocks.1." + n].contiguous() for n in ops.NSTB_PARAM_NAMES]
B, ph, pw = 32, 416, 416
z = torch.randn(B, ph * pw, 64, device="cuda")
c = torch.randn(B, (ph // 8) * (pw // 8), 64, device="cuda") * 0.5
buf = (ctypes.c_ulonglong * 32)()
ops.attn_tc_fwd(z, c, *P[10:18], ph, pw, 6, 0); torch.cuda.synchronize()
L.b200mar_attn_prof(buf, 1)
ops.attn_tc_fwd(z, c, *P[10:18], ph, pw, 6, 0); torch.cuda.synchronize()
L.b200mar_attn_prof(buf, 1)
names = ["loop top", "ctx + wait z tile", "sync+qkv0 issue", "wait qkv", "qkv ld+norm+pack", "wait PV+O read", "operand stores",
         "fence+sync", "S issue+wait", "softmax+P st", "sync(or)", "PV issue", "Otile+proj+wait", "LN epilogue", "wait z reload",
         "a in place+store+next load", "PV issue (last)", "row convert+mask"]
tot = sum(buf)
ntile = (B * (ph // 8) * (pw // 8) // 2 + 148 * 4 - 1) // (148 * 4)   # 4 workers per CTA (NH = 6)
print("tiles per worker ~", ntile, "total cycles", tot, "per tile", tot / ntile)
for i, n in enumerate(names):
    print(f"{i:2d} {n:22s} {buf[i] / ntile:9.0f} cyc/tile  {100 * buf[i] / tot:5.1f}%")
