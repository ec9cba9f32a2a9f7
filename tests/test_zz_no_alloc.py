"""b200mar_ngram_ctx_infer_noalloc (no counterpart in the reference):
the n-gram context inference pair with its nq scratch in a static __device__ array instead of a caller-allocated
buffer.  Called straight through the C ABI (ctypes) and compared with the existing two-call entry points on two
calls in a row (the shared scratch keeps the previous call's contents: the second call uses a different shape and
different data).  Same kernels, same launch geometry -> the results must be bit-identical (tolerance 0).
Named test_zz_* so that pytest collects it last."""
import pytest
import torch

from oracle import ngswin_oracle as O

pytestmark = pytest.mark.gpu

NGRAM = ("unigram_embed.weight", "unigram_embed.bias", "ngram_attn.qkv.weight", "ngram_attn.qkv.bias",
         "ngram_attn.proj.weight", "ngram_attn.proj.bias", "ngram_attn.logit_scale",
         "ngram_attn.relative_position_bias_table", "merge.weight", "merge.bias")


def _params(prefix, kind):
    sd = O.synthetic_state_dict(kind, seed=1)
    pre = prefix + "ngram_window_partition.ngram_context."
    return [sd[pre + n].cuda().contiguous() for n in NGRAM]


def _two_call(lib, z, P, B, ph, pw, nh, stream):
    wh, ww = ph // 8, pw // 8
    nq = torch.empty(B * wh * ww, 96, device="cuda")
    ctx = torch.full((B, wh, ww, 64), float("nan"), device="cuda")
    p = [t.data_ptr() for t in P]
    assert lib.b200mar_unigram_qkv_fwd(z.data_ptr(), p[0], p[1], p[2], p[3], nq.data_ptr(), B, ph, pw, nh, stream) == 0
    assert lib.b200mar_ngram_ctx_pre_fwd(nq.data_ptr(), p[4], p[5], p[6], p[7], p[8], p[9], ctx.data_ptr(),
                                         B, wh, ww, nh, stream) == 0
    return ctx


def _no_alloc(lib, z, P, ctx, B, ph, pw, nh, stream):
    p = [t.data_ptr() for t in P]
    return lib.b200mar_ngram_ctx_infer_noalloc(z.data_ptr(), *p, ctx.data_ptr(), B, ph, pw, nh, stream)


def test_no_alloc_entry_point_matches_two_call_form_twice_in_a_row():
    from transmar_gan_b200 import _cabi
    assert torch.cuda.is_available(), "gpu tests need a CUDA device"
    lib = _cabi.lib()
    assert lib.b200mar_device_ok(0) == 1
    assert lib.b200mar_ngram_ctx_infer_noalloc_max_windows() == 16384
    stream = torch.cuda.current_stream().cuda_stream
    g = torch.Generator().manual_seed(17)
    cases = [("encoder_layer1.blocks.1.", "W2", 6, 2, 96, 160),      # call 1: 480 windows, 6 heads
             ("encoder_layer2.blocks.0.", "W1", 4, 3, 48, 32)]       # call 2: 72 windows, 4 heads, other data
    for prefix, kind, nh, B, ph, pw in cases:
        z = torch.randn(B, ph * pw, 64, generator=g).cuda()
        P = _params(prefix, kind)
        ref = _two_call(lib, z, P, B, ph, pw, nh, stream)
        ctx = torch.full_like(ref, float("nan"))                     # outputs are allocated BEFORE the call
        torch.cuda.synchronize()
        before = torch.cuda.memory_stats()["allocation.all.allocated"]
        rc = _no_alloc(lib, z, P, ctx, B, ph, pw, nh, stream)
        after = torch.cuda.memory_stats()["allocation.all.allocated"]
        torch.cuda.synchronize()
        assert rc == 0, _cabi.lib().b200mar_last_error()
        assert after == before                                       # torch's allocator was not touched either
        assert torch.isfinite(ctx).all()
        assert torch.equal(ctx, ref), (ctx - ref).abs().max().item()
        # and against the oracle (the tolerance tests/test_gpu_parity.py uses for ctx)
        sd = O.synthetic_state_dict(kind, seed=1)
        with torch.no_grad():
            want = O.ngram_context(z.cpu().view(B, ph, pw, 64), sd,
                                   prefix + "ngram_window_partition.ngram_context.", nh)
        assert (ctx.cpu().reshape(want.shape) - want).abs().max().item() < 3e-4


def test_no_alloc_entry_point_rejects_oversize_and_bad_geometry():
    from transmar_gan_b200 import _cabi
    lib = _cabi.lib()
    stream = torch.cuda.current_stream().cuda_stream
    P = _params("encoder_layer1.blocks.0.", "W1")
    z = torch.zeros(1, 64, device="cuda")                            # never read: the call must fail before launching
    ctx = torch.zeros(1, 64, device="cuda")
    rc = _no_alloc(lib, z, P, ctx, 2, 1024, 1024, 6, stream)         # 32768 windows > 16384
    assert rc == -5
    assert b"16384" in lib.b200mar_last_error()
    assert _no_alloc(lib, z, P, ctx, 1, 8, 64, 6, stream) == -1      # a single window row: the reference raises too
    assert _no_alloc(lib, z, P, ctx, 1, 64, 64, 5, stream) == -1
    torch.cuda.synchronize()
