"""Development aid: device time of unigram_bwd (+dz_add) at the training shape (batch 16, 128x128 tokens)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from transmar_gan_b200 import ops
B, S = 16, 128
z = torch.randn(B, S * S, 64, device="cuda"); w = torch.randn(32, 2, 8, 8, device="cuda")
du = torch.randn(B, S // 8, S // 8, 32, device="cuda"); dz = torch.randn_like(z); add = torch.randn_like(z)
dw = torch.zeros_like(w); db = torch.zeros(32, device="cuda")
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
for mult in ["library default"]:
    ts = []
    for it in range(12):
        flush.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); ops.unigram_bwd(z, w, du, dz, dw, db, S, S, add); e1.record()
        torch.cuda.synchronize(); ts.append(e0.elapsed_time(e1) * 1e3)
    ts = sorted(ts[2:])
    print(f"grid = {mult}: median {ts[len(ts)//2]:.1f} us, min {ts[0]:.1f} us (L2 flushed before each launch)")
