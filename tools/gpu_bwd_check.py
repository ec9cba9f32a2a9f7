"""Scratch GPU check of the tensor-core backward kernels against the fp32 SIMT backward kernels."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import transmar_gan_b200 as T
from oracle import ngswin_oracle as O

ops = T.ops
torch.manual_seed(0)


def timeit(fn, n=10):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n


def rel(x, y):
    return ((x - y).norm() / y.norm().clamp_min(1e-30)).item()


# ---- wgrad alone
for (T_, cin, cout) in ((128, 64, 128), (192, 64, 192), (4096 + 64, 128, 64), (100032, 64, 192)):
    a = torch.randn(T_, cin, device="cuda").bfloat16()
    g = torch.randn(T_, cout, device="cuda").bfloat16()
    dw = torch.zeros(cout, cin, device="cuda")
    db = torch.zeros(cout, device="cuda") if cin == 64 else None
    ops.wgrad_tc(a, g, dw, db)
    ref = g.double().T @ a.double()
    msg = f"wgrad T={T_} cin={cin} cout={cout}: dw rel {rel(dw.double(), ref):.2e}"
    if db is not None:
        msg += f" dbias rel {rel(db.double(), g.double().sum(0)):.2e}"
    print(msg, "tc_error", ops.tc_error(), flush=True)

# ---- MLP backward
for kind in ("W1", "W2"):
    sd = {k: v.cuda() for k, v in O.synthetic_state_dict(kind, seed=1).items()}
    P = [sd["encoder_layer1.blocks.1." + n].contiguous() for n in ops.NSTB_PARAM_NAMES][18:24]
    for n_tok in (64, 128, 4096 + 64, 16 * 128 * 128):
        a = torch.randn(n_tok, 64, device="cuda")
        g = torch.randn(n_tok, 64, device="cuda") * 1e-3
        outs = []
        for fn in (ops.mlp_bwd, ops.mlp_bwd_tc):
            da = torch.empty_like(a)
            G = [torch.zeros_like(p) for p in P]
            fn(a, g, P, da, G)
            outs.append((da, G))
        torch.cuda.synchronize()
        (da0, G0), (da1, G1) = outs
        names = ("fc1_w", "fc1_b", "fc2_w", "fc2_b", "ln_w", "ln_b")
        print(f"{kind} mlp_bwd n_tok={n_tok}: da rel {rel(da1, da0):.2e} | " +
              " ".join(f"{n} {rel(x, y):.2e}" for n, x, y in zip(names, G1, G0)),
              "tc_error", ops.tc_error(), flush=True)
    n_tok = 16 * 128 * 128
    a = torch.randn(n_tok, 64, device="cuda")
    g = torch.randn(n_tok, 64, device="cuda")
    da = torch.empty_like(a)
    G = [torch.zeros_like(p) for p in P]
    print(f"mlp_bwd simt {timeit(lambda: ops.mlp_bwd(a, g, P, da, G)):.3f} ms, "
          f"tc {timeit(lambda: ops.mlp_bwd_tc(a, g, P, da, G)):.3f} ms (n_tok={n_tok})", flush=True)

# ---- attention backward
names = ("qkv_w", "qkv_b", "proj_w", "proj_b", "logit_scale", "bias_table", "ln_w", "ln_b")
for kind in ("W1", "W2"):
    sd = {k: v.cuda() for k, v in O.synthetic_state_dict(kind, seed=1).items()}
    for pre, nh in (("encoder_layer1.blocks.1.", 6), ("encoder_layer2.blocks.1.", 4)):
        P = [sd[pre + n].contiguous() for n in ops.NSTB_PARAM_NAMES][10:18]
        for (B, ph, pw) in ((1, 16, 16), (1, 24, 40), (3, 24, 24)):
            for shift in (0, 4):
                z = torch.randn(B, ph * pw, 64, device="cuda")
                c = torch.randn(B, (ph // 8) * (pw // 8), 64, device="cuda") * 0.5
                da = torch.randn_like(z) * 1e-2
                outs = []
                for fn in (ops.attn_bwd, ops.attn_bwd_tc):
                    dz = torch.empty_like(z)
                    dctx = torch.zeros_like(c)
                    G = [torch.zeros_like(p) for p in P]
                    fn(z, c, da, P, dz, dctx, G, ph, pw, nh, shift)
                    outs.append((dz, dctx, G))
                torch.cuda.synchronize()
                a_s, o16 = ops.attn_tc_fwd_save(z, c, *P, ph, pw, nh, shift)
                dz = torch.empty_like(z); dctx = torch.zeros_like(c); G = [torch.zeros_like(p) for p in P]
                ops.attn_bwd_tc(z, c, da, P, dz, dctx, G, ph, pw, nh, shift, o16)
                (dz0, dc0, G0), (dz1, dc1, G1) = outs
                a_ref = ops.attn_tc_fwd(z, c, *P, ph, pw, nh, shift)
                print(f"   saved-O path: fwd_save vs fwd max {(a_s - a_ref).abs().max().item():.1e}; dz-da rel {rel(dz - da, dz0 - da):.2e} "
                      f"dctx {rel(dctx, dc0):.2e} | " + " ".join(f"{n} {rel(x, y):.2e}" for n, x, y in zip(names, G, G0)))
                print(f"{kind} attn_bwd nh={nh} B={B} {ph}x{pw} shift={shift}: dz-da rel {rel(dz1 - da, dz0 - da):.2e} "
                      f"dctx {rel(dc1, dc0):.2e} | " + " ".join(f"{n} {rel(x, y):.2e}" for n, x, y in zip(names, G1, G0)),
                      "tc_error", ops.tc_error(), flush=True)
sd = {k: v.cuda() for k, v in O.synthetic_state_dict("W1", seed=1).items()}
for pre, nh, (B, ph, pw) in (("encoder_layer1.blocks.1.", 6, (16, 128, 128)), ("encoder_layer2.blocks.1.", 4, (16, 64, 64))):
    P = [sd[pre + n].contiguous() for n in ops.NSTB_PARAM_NAMES][10:18]
    z = torch.randn(B, ph * pw, 64, device="cuda")
    c = torch.randn(B, (ph // 8) * (pw // 8), 64, device="cuda") * 0.5
    da = torch.randn_like(z)
    dz = torch.empty_like(z)
    dctx = torch.zeros_like(c)
    G = [torch.zeros_like(p) for p in P]
    for shift in (0, 4):
        print(f"attn_bwd nh={nh} shift={shift} B={B} {ph}x{pw}: simt "
              f"{timeit(lambda: ops.attn_bwd(z, c, da, P, dz, dctx, G, ph, pw, nh, shift)):.3f} ms, tc "
              f"{timeit(lambda: ops.attn_bwd_tc(z, c, da, P, dz, dctx, G, ph, pw, nh, shift)):.3f} ms, tc with saved O "
              f"{timeit(lambda: ops.attn_bwd_tc(z, c, da, P, dz, dctx, G, ph, pw, nh, shift, ops.attn_tc_fwd_save(z, c, *P, ph, pw, nh, shift)[1])):.3f} ms (incl. fwd), fwd_save "
              f"{timeit(lambda: ops.attn_tc_fwd_save(z, c, *P, ph, pw, nh, shift)):.3f} ms", flush=True)
