"""Development aid: run-to-run noise of the backward vs the fused-accumulation difference, per parameter."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch, torch.nn.functional as F
import transmar_gan_b200 as T
from transmar_gan_b200 import dp, ops
from oracle import ngswin_oracle as O
m = T.NGswin(); m.load_state_dict(O.synthetic_state_dict("W1", seed=1), strict=True); m = m.cuda().train()
bucket = dp.FlatGradBucket(m.parameters())
x = O.synthetic_input(2, 64, 96, seed=999).cuda(); t = O.synthetic_input(2, 64, 96, seed=1000).cuda()
def run(fused):
    bucket.zero()
    loss = F.l1_loss(m(x), t)
    with ops.fused_grad_accumulation(fused):
        loss.backward()
    return {n: p.grad.clone() for n, p in m.named_parameters()}
a, b, c, d = run(False), run(False), run(True), run(True)
def report(u, v, tag):
    tot = sum((u[n] - v[n]).norm() ** 2 for n in u).sqrt(); ref = sum(u[n].norm() ** 2 for n in u).sqrt()
    worst = sorted(((float((u[n] - v[n]).norm() / (u[n].norm() + 1e-20)), n) for n in u), reverse=True)[:6]
    print(tag, float(tot / ref), worst)
report(a, b, "unfused vs unfused")
report(c, d, "fused vs fused")
report(a, c, "unfused vs fused")
