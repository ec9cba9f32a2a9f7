"""Development aid: mlp_bwd_tc (row kernel + 2 wgrad launches) with the TMA row kernel vs the plain-load one
(B200MAR_MLP_BWD_PLAIN=1 in the environment selects the latter): results and time at the training shape."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import transmar_gan_b200 as T
from oracle import ngswin_oracle as O
ops = T.ops
sd = {k: v.cuda() for k, v in O.synthetic_state_dict("W1", seed=1).items()}
P = [sd["encoder_layer1.blocks.1." + n].contiguous() for n in ops.NSTB_PARAM_NAMES][18:24]
for n_tok in (16 * 128 * 128, 64 * 37, 128 * 5 + 64):
    g = torch.Generator(device="cuda").manual_seed(n_tok)
    a = torch.randn(1, n_tok, 64, device="cuda", generator=g); dout = torch.randn(1, n_tok, 64, device="cuda", generator=g)
    da = torch.empty_like(a); G = [torch.zeros_like(p) for p in P]
    ops.mlp_bwd_tc(a, dout, P, da, G); torch.cuda.synchronize()
    print(n_tok, "da sum", float(da.double().sum()), "abs", float(da.abs().double().sum()), "grads", [round(float(x.double().abs().sum()), 4) for x in G])
    if n_tok > 100000:
        for _ in range(3): ops.mlp_bwd_tc(a, dout, P, da, G)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(20): ops.mlp_bwd_tc(a, dout, P, da, G)
        e1.record(); torch.cuda.synchronize()
        print("mlp_bwd_tc (row kernel + 2 wgrad)", round(e0.elapsed_time(e1) / 20 * 1e3, 1), "us")
print("tc_error", ops.tc_error())
