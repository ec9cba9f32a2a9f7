"""Development aid: per-phase cycle profile of mlp_tc (library built with -DB200MAR_MLP_PROF)."""
import sys, os, ctypes
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import transmar_gan_b200 as T
from oracle import ngswin_oracle as O
ops = T.ops
L = ctypes.CDLL(T.LIB_PATH)
L.b200mar_mlp_prof.argtypes = [ctypes.c_void_p, ctypes.c_int]
sd = {k: v.cuda() for k, v in O.synthetic_state_dict("W1", seed=1).items()}
P = [sd["encoder_layer1.blocks.1." + n].contiguous() for n in ops.NSTB_PARAM_NAMES]
B, T_ = 32, 416 * 416
a = torch.randn(B, T_, 64, device="cuda")
z = torch.randn(B, T_, 64, device="cuda")
buf = (ctypes.c_ulonglong * 32)()
for _ in range(2):
    ops.mlp_tc_fwd(a, z, *P[18:24], False, True); torch.cuda.synchronize()
    L.b200mar_mlp_prof(buf, 1)
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(5): ops.mlp_tc_fwd(a, z, *P[18:24], False, True)
e1.record(); torch.cuda.synchronize()
print("ms per launch", e0.elapsed_time(e1) / 5)
L.b200mar_mlp_prof(buf, 1)
names = ["loop top", "load+convert", "sync", "fc1 issue+wait", "GELU", "sync", "fc2 issue+wait", "LN", "sync", "copy-out", "sync"]
tot = sum(buf) / 5
ntile = (B * T_ // 128 + 148 * 4 - 1) // (148 * 4)
print("tiles per worker ~", ntile, "per tile", tot / ntile)
for i, n in enumerate(names):
    print(f"{i:2d} {n:16s} {buf[i] / 5 / ntile:9.0f} cyc/tile  {100 * buf[i] / 5 / tot:5.1f}%")
