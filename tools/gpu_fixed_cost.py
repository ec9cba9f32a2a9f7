"""Development aid: fixed cost of one launch of each block kernel (prologue + one tile + exit), measured at a
tiny shape (1 x 16 x 16 tokens = 4 windows = 2 tiles) as 50 back-to-back launches inside one CUDA graph."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import transmar_gan_b200 as T
from oracle import ngswin_oracle as O
ops = T.ops
sd = {k: v.cuda() for k, v in O.synthetic_state_dict("W1", seed=1).items()}
B, ph, pw = 1, 16, 16
N = 50


def bench(name, f):
    f(); torch.cuda.synchronize()
    s = torch.cuda.Stream()
    s.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(s):
        f(); f()
    torch.cuda.current_stream().wait_stream(s)
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        for _ in range(N):
            f()
    g.replay(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(); g.replay(); g.replay(); e1.record(); torch.cuda.synchronize()
    print(f"{name:28s} {e0.elapsed_time(e1) / (2 * N) * 1e3:7.1f} us per call")


for pre, nh in (("encoder_layer1.blocks.1.", 6), ("encoder_layer2.blocks.1.", 4)):
    P = [sd[pre + n].contiguous() for n in ops.NSTB_PARAM_NAMES]
    z = torch.randn(B, ph * pw, 64, device="cuda")
    da = torch.randn_like(z); dz = torch.zeros_like(z)
    G = [torch.zeros_like(p) for p in P]
    u = ops.unigram_fwd(z, P[0], P[1], ph, pw)
    ctx = ops.ngram_ctx_fwd(u, *P[2:10], nh)
    dctx = torch.zeros_like(ctx); du = torch.zeros_like(u)
    a, o16 = ops.attn_tc_fwd_save(z, ctx.view(B, -1, 64), *P[10:18], ph, pw, nh, 4)
    c = ctx.view(B, -1, 64)
    print(f"--- {nh} heads")
    bench("unigram_fwd", lambda: ops.unigram_fwd(z, P[0], P[1], ph, pw))
    bench("ngram_ctx_fwd", lambda: ops.ngram_ctx_fwd(u, *P[2:10], nh))
    bench("ngram_ctx_infer (2 kernels)", lambda: ops.ngram_ctx_infer(z, *P[0:10], ph, pw, nh))
    bench("attn_tc_fwd", lambda: ops.attn_tc_fwd(z, c, *P[10:18], ph, pw, nh, 4))
    bench("attn_tc_fwd_save", lambda: ops.attn_tc_fwd_save(z, c, *P[10:18], ph, pw, nh, 4))
    bench("mlp_tma_fwd", lambda: ops.mlp_tma_fwd_(a, z, *P[18:24], False))
    bench("attn_bwd_tc (4 kernels)", lambda: ops.attn_bwd_tc(z, c, da, P[10:18], dz, dctx.view(B, -1, 64), G[10:18], ph, pw, nh, 4, o16))
    bench("mlp_bwd_tc (3 kernels)", lambda: ops.mlp_bwd_tc(a.view(-1, 64), da.view(-1, 64), P[18:24], dz.view(-1, 64), G[18:24]))
    bench("ngram_ctx_bwd (+memset)", lambda: ops.ngram_ctx_bwd(u, dctx, P[2:10], du, G[2:10], nh))
    bench("unigram_bwd", lambda: ops.unigram_bwd(z, P[0], du, dz, G[0], G[1], ph, pw))
x16 = torch.randn(256, 64, device="cuda").bfloat16(); g16 = torch.randn(256, 192, device="cuda").bfloat16()
dw = torch.zeros(192, 64, device="cuda"); db = torch.zeros(192, device="cuda")
bench("wgrad_tc 64->192", lambda: ops.wgrad_tc(x16, g16, dw, db))
e = torch.empty(1, device="cuda")
bench("empty torch kernel (fill)", lambda: e.fill_(1.0))
print("tc_error", ops.tc_error())
