"""Scratch GPU check of the tensor-core kernels against the fp32 SIMT kernels (development aid)."""
import sys, os, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import transmar_gan_b200 as T
from oracle import ngswin_oracle as O

ops = T.ops
torch.manual_seed(0)


def timeit(fn, n=5):
    for _ in range(2):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n


for kind in ("W1", "W2"):
    sd = {k: v.cuda() for k, v in O.synthetic_state_dict(kind, seed=1).items()}
    for pre, nh in (("encoder_layer1.blocks.1.", 6), ("encoder_layer2.blocks.1.", 4)):
        P = [sd[pre + n].contiguous() for n in ops.NSTB_PARAM_NAMES]
        for (B, ph, pw) in ((1, 16, 16), (1, 24, 40), (3, 24, 24), (2, 64, 64)):
            for shift in (0, 4):
                z = torch.randn(B, ph * pw, 64, device="cuda")
                c = torch.randn(B, (ph // 8) * (pw // 8), 64, device="cuda") * 0.5
                a1 = ops.attn_fwd(z, c, *P[10:18], ph, pw, nh, shift)
                a2 = ops.attn_tc_fwd(z, c, *P[10:18], ph, pw, nh, shift)
                torch.cuda.synchronize()
                d = (a1 - a2).abs()
                print(f"{kind} nh={nh} B={B} {ph}x{pw} shift={shift}: attn tc vs simt max {d.max().item():.3e} "
                      f"mean {d.mean().item():.3e} (|a-z| max {(a1 - z).abs().max().item():.2f}) "
                      f"tc_error {ops.tc_error()}", flush=True)
sd = {k: v.cuda() for k, v in O.synthetic_state_dict("W1", seed=1).items()}
for pre, nh, (B, ph, pw) in (("encoder_layer1.blocks.1.", 6, (32, 416, 416)), ("encoder_layer2.blocks.1.", 4, (32, 208, 208))):
    P = [sd[pre + n].contiguous() for n in ops.NSTB_PARAM_NAMES]
    z = torch.randn(B, ph * pw, 64, device="cuda")
    c = torch.randn(B, (ph // 8) * (pw // 8), 64, device="cuda") * 0.5
    for shift in (0, 4):
        ms = timeit(lambda: ops.attn_tc_fwd(z, c, *P[10:18], ph, pw, nh, shift))
        gb = z.numel() * 4 * 2 / 1e9
        print(f"attn tc nh={nh} shift={shift} B={B} {ph}x{pw}: {ms:.3f} ms per launch, {gb / ms * 1e3:.0f} GB/s "
              f"algorithmic (2 passes), tc_error {ops.tc_error()}", flush=True)
    a = torch.randn_like(z)
    ms = timeit(lambda: ops.mlp_tc_fwd(a, z, *P[18:24], False, True, True, False))
    for shift in (0, 4):
        ms2 = timeit(lambda: ops.attn_tc_fwd(z, c, *P[10:18], ph, pw, nh, shift, False))
        print(f"attn tc (no residual) nh={nh} shift={shift}: {ms2:.3f} ms per launch")
    print(f"mlp tc B={B} {ph}x{pw}: {ms:.3f} ms per launch, {z.numel() * 4 * 3 / 1e9 / ms * 1e3:.0f} GB/s algorithmic (3 passes)")
