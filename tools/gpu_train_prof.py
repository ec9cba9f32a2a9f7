"""Development aid: a few config-3 training steps (fwd + L1 + bwd + Adam, batch 16 x 128x128) through
train.L1TrainStep with eager launches, for ncu launch lists (same recipe as bench.py's train leg)."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import bench
from transmar_gan_b200 import train
torch.backends.cuda.matmul.allow_tf32 = False
torch.backends.cudnn.allow_tf32 = True   # as bench.py's train leg (torch default)
torch.backends.cudnn.benchmark = True
m = bench.build_model().train()
step = train.L1TrainStep(m, lr=1e-4, betas=(0.5, 0.999), graph=False)
B = int(sys.argv[1]) if len(sys.argv) > 1 else 16
x = (torch.rand(B, 1, 128, 128) * 2 - 1).cuda(); t = (torch.rand(B, 1, 128, 128) * 2 - 1).cuda()
for i in range(4):
    step(x, t)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
torch.cuda.profiler.start()            # ncu --profile-from-start off: only the timed steps are listed
e0.record()
for i in range(3):
    loss = step(x, t)
e1.record(); torch.cuda.synchronize()
torch.cuda.profiler.stop()
print("ms/step", e0.elapsed_time(e1) / 3, "loss", loss.item())
