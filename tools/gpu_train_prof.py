"""Development aid: a few config-3 training steps (fwd + L1 + bwd + Adam, batch 16 x 128x128) for ncu."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch, torch.nn.functional as F
import bench
torch.backends.cuda.matmul.allow_tf32 = False
torch.backends.cudnn.allow_tf32 = False
m = bench.build_model().train()
opt = torch.optim.Adam(m.parameters(), lr=1e-4, betas=(0.5, 0.999))
B = int(sys.argv[1]) if len(sys.argv) > 1 else 16
x = (torch.rand(B, 1, 128, 128) * 2 - 1).cuda(); t = (torch.rand(B, 1, 128, 128) * 2 - 1).cuda()
for i in range(4):
    opt.zero_grad(set_to_none=True)
    loss = F.l1_loss(m(x), t); loss.backward(); opt.step()
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for i in range(3):
    opt.zero_grad(set_to_none=True)
    loss = F.l1_loss(m(x), t); loss.backward(); opt.step()
e1.record(); torch.cuda.synchronize()
print("ms/step", e0.elapsed_time(e1) / 3, "loss", loss.item())
