"""Development aid: CUDA-event timing of ngram_ctx_bwd / unigram_bwd at the training size (16 x 128^2)."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import transmar_gan_b200 as T
from oracle import ngswin_oracle as O
ops = T.ops
sd = {k: v.cuda() for k, v in O.synthetic_state_dict("W1", seed=1).items()}
B, ph, pw = 16, 128, 128
P = [sd["encoder_layer1.blocks.1." + n].contiguous() for n in ops.NSTB_PARAM_NAMES]
z = torch.randn(B, ph * pw, 64, device="cuda"); dz = torch.zeros_like(z)
G = [torch.zeros_like(p) for p in P]
u = ops.unigram_fwd(z, P[0], P[1], ph, pw)
ctx = ops.ngram_ctx_fwd(u, *P[2:10], 6)
dctx = torch.randn_like(ctx); du = torch.zeros_like(u)


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


a = t(lambda: ops.ngram_ctx_bwd(u, dctx, P[2:10], du, G[2:10], 6))
b = t(lambda: ops.unigram_bwd(z, P[0], du, dz, G[0], G[1], ph, pw))
c = t(lambda: ops.ngram_ctx_fwd(u, *P[2:10], 6))
d = t(lambda: ops.unigram_fwd(z, P[0], P[1], ph, pw))
e = t(lambda: ops.ngram_ctx_infer(z, *P[0:10], ph, pw, 6))
print(f"{os.path.basename(T.LIB_PATH)} | ngram_ctx_bwd {a:.1f} us | unigram_bwd {b:.1f} us | ngram_ctx_fwd {c:.1f} us | unigram_fwd {d:.1f} us | ngram_ctx_infer (unigram_qkv + ctx_pre) {e:.1f} us")
