"""Development aid: which Python frames launch the small fill kernels of the training step."""
import sys, os, collections
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch, torch.nn.functional as F
import bench
from transmar_gan_b200 import train
from torch.profiler import profile, ProfilerActivity
m = bench.build_model().train()
step = train.L1TrainStep(m, graph=False)
x = (torch.rand(16, 1, 128, 128) * 2 - 1).cuda(); t = (torch.rand(16, 1, 128, 128) * 2 - 1).cuda()
for _ in range(3):
    step(x, t)
torch.cuda.synchronize()
with profile(activities=[ProfilerActivity.CPU, ProfilerActivity.CUDA], with_stack=True) as prof:
    step(x, t)
    torch.cuda.synchronize()
c = collections.Counter()
for e in prof.events():
    if e.name in ("aten::fill_", "aten::zero_", "aten::zeros", "aten::zeros_like", "aten::add_", "aten::add"):
        st = [s for s in (e.stack or []) if "site-packages/torch" not in s][:3]
        c[(e.name, tuple(st))] += 1
for (n, st), k in c.most_common(25):
    print(k, n, " | ".join(s[-70:] for s in st))
