"""Development aid: mlp_tma_fwd_ vs mlp_tc_fwd (same arithmetic) + timing at B=32 x 416^2."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import transmar_gan_b200 as T
from oracle import ngswin_oracle as O
ops = T.ops
sd = {k: v.cuda() for k, v in O.synthetic_state_dict("W1", seed=1).items()}
P = [sd["encoder_layer1.blocks.1." + n].contiguous() for n in ops.NSTB_PARAM_NAMES]
W = P[18:24]
torch.manual_seed(0)
for ntok in (64, 128, 192, 4096 + 64, 148 * 3 * 128 * 2 + 64):
    a = torch.randn(1, ntok, 64, device="cuda") * 3
    z = torch.randn(1, ntok, 64, device="cuda")
    a[0, 0] = 0; a[0, 1] *= 1e-6; a[0, 2] *= 1e5
    o_ref, zs_ref, _ = ops.mlp_tc_fwd(a, z, *W, True, True)
    z2 = z.clone()
    o = ops.mlp_tma_fwd_(a, z2, *W, True)
    torch.cuda.synchronize()
    print(ntok, "out max diff", (o - o_ref).abs().max().item(), "zsum max diff", (z2 - zs_ref).abs().max().item(),
          "tc_error", ops.tc_error())
    z3 = z.clone()
    ops.mlp_tma_fwd_(a, z3, *W, False)
    o4 = ops.mlp_tma_fwd_(a, None, *W, True)
    print("   zsum-only diff", (z3 - zs_ref).abs().max().item(), "out-only diff", (o4 - o_ref).abs().max().item())
B, T_ = 32, 416 * 416
a = torch.randn(B, T_, 64, device="cuda")
z = torch.randn(B, T_, 64, device="cuda")
def timeit(f, n=5):
    f(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n): f()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n
print("mlp_tc_fwd  (zsum)      ms", timeit(lambda: ops.mlp_tc_fwd(a, z, *W, False, True)))
print("mlp_tma_fwd_ (z += out) ms", timeit(lambda: ops.mlp_tma_fwd_(a, z, *W, False)))
print("mlp_tma_fwd_ (out only) ms", timeit(lambda: ops.mlp_tma_fwd_(a, None, *W, True)))
print("mlp_tma_fwd_ (both)     ms", timeit(lambda: ops.mlp_tma_fwd_(a, z, *W, True)))
print("tc_error", ops.tc_error())
