"""Development aid: which backward op is not repeatable? One NSTB, ops called directly, repeated."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import transmar_gan_b200 as T
from oracle import ngswin_oracle as O
ops = T.ops
sd = {k: v.cuda() for k, v in O.synthetic_state_dict("W1", seed=1).items()}
def rel(a, b): return float((a - b).norm() / (b.norm() + 1e-30))
for prefix, nh, shift, B, ph, pw in [("encoder_layer1.blocks.1.", 6, 4, 2, 64, 96), ("encoder_layer1.blocks.0.", 6, 0, 2, 64, 96),
                                     ("encoder_layer2.blocks.1.", 4, 4, 2, 32, 48)]:
    P = [sd[prefix + n].contiguous() for n in ops.NSTB_PARAM_NAMES]
    g = torch.Generator().manual_seed(3)
    z = torch.randn(B, ph * pw, 64, generator=g).cuda()
    dout = torch.randn(B, ph * pw, 64, generator=g).cuda()
    u = ops.unigram_fwd(z, P[0], P[1], ph, pw)
    c = ops.ngram_ctx_fwd(u, *P[2:10], nh)
    a, o16 = ops.attn_tc_fwd_save(z, c, *P[10:18], ph, pw, nh, shift)
    res = {k: [] for k in ("mlp da", "mlp G", "attn dz", "attn dctx", "attn G", "attn(no o16) dz", "ngram du", "ngram G", "uni dz", "uni G")}
    for it in range(5):
        G = [torch.zeros_like(p) for p in P]
        da = torch.empty_like(a)
        ops.mlp_bwd_tc(a, dout, P[18:24], da, G[18:24])
        res["mlp da"].append(da.clone()); res["mlp G"].append(torch.cat([t.flatten() for t in G[18:24]]))
        da0 = res["mlp da"][0]
        dz = torch.empty_like(z); dctx = torch.zeros_like(c)
        ops.attn_bwd_tc(z, c, da0, P[10:18], dz, dctx, G[10:18], ph, pw, nh, shift, o16)
        res["attn dz"].append(dz.clone()); res["attn dctx"].append(dctx.clone()); res["attn G"].append(torch.cat([t.flatten() for t in G[10:18]]))
        G2 = [torch.zeros_like(p) for p in P]
        dz2 = torch.empty_like(z); dctx2 = torch.zeros_like(c)
        ops.attn_bwd_tc(z, c, da0, P[10:18], dz2, dctx2, G2[10:18], ph, pw, nh, shift, None)
        res["attn(no o16) dz"].append(dz2.clone())
        dctx0 = res["attn dctx"][0]
        du = torch.empty_like(u)
        ops.ngram_ctx_bwd(u, dctx0, P[2:10], du, G[2:10], nh)
        res["ngram du"].append(du.clone()); res["ngram G"].append(torch.cat([t.flatten() for t in G[2:10]]))
        dzz = res["attn dz"][0].clone()
        ops.unigram_bwd(z, P[0], res["ngram du"][0], dzz, G[0], G[1], ph, pw)
        res["uni dz"].append(dzz); res["uni G"].append(torch.cat([G[0].flatten(), G[1].flatten()]))
    print(prefix, "shift", shift)
    for k, v in res.items():
        print(f"  {k:18s}", ["%.1e" % rel(t, v[0]) for t in v[1:]])
print("tc_error", ops.tc_error())
