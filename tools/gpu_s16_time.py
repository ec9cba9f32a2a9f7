"""Development aid: per-launch time of the three per-block kernels, fp32 stream vs bf16 stream, batch 32 x 416^2."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import transmar_gan_b200 as T
from oracle import ngswin_oracle as O
ops, lib = T.ops, T._cabi.lib()
B, ph, pw, nh = 32, 416, 416, 6
sd = {k: v.cuda() for k, v in O.synthetic_state_dict("W1", seed=1).items()}
P = [sd["encoder_layer1.blocks.1." + n].contiguous() for n in ops.NSTB_PARAM_NAMES]
p = [t.data_ptr() for t in P]
z = torch.randn(B, ph * pw, 64, device="cuda"); z16 = z.bfloat16()
a = torch.empty_like(z); a16 = torch.empty_like(z16)
ctx = torch.randn(B, (ph // 8) * (pw // 8), 64, device="cuda") * 0.5
nq = torch.empty(B, ph // 8, pw // 8, 96, device="cuda")
st = torch.cuda.current_stream().cuda_stream
n = B * ph * pw
def timeit(name, fn, reps=10):
    for _ in range(3): fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps): fn()
    e1.record(); torch.cuda.synchronize()
    print(f"{name:34s} {e0.elapsed_time(e1) / reps:8.4f} ms", flush=True)
for shift in (0, 4):
    timeit(f"attn_tc_fwd      shift {shift}", lambda: lib.b200mar_attn_tc_fwd(z.data_ptr(), ctx.data_ptr(), *p[10:18], a.data_ptr(), B, ph, pw, nh, shift, 1, st))
    timeit(f"attn_tc_fwd_s16  shift {shift}", lambda: lib.b200mar_attn_tc_fwd_s16(z16.data_ptr(), ctx.data_ptr(), *p[10:18], a16.data_ptr(), B, ph, pw, nh, shift, st))
timeit("mlp_tma_fwd      (z += out)", lambda: lib.b200mar_mlp_tma_fwd(a.data_ptr(), z.data_ptr(), *p[18:24], None, n, st))
timeit("mlp_tma_fwd_s16  (z += out)", lambda: lib.b200mar_mlp_tma_fwd_s16(a16.data_ptr(), z16.data_ptr(), *p[18:24], None, n, st))
timeit("unigram_qkv_fwd", lambda: lib.b200mar_unigram_qkv_fwd(z.data_ptr(), *p[0:4], nq.data_ptr(), B, ph, pw, nh, st))
timeit("unigram_qkv_fwd_s16", lambda: lib.b200mar_unigram_qkv_fwd_s16(z16.data_ptr(), *p[0:4], nq.data_ptr(), B, ph, pw, nh, st))
timeit("cast fp32 -> bf16", lambda: z.to(torch.bfloat16))
timeit("cast bf16 -> fp32", lambda: z16.float())
print("tc_error", ops.tc_error())
