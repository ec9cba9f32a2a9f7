"""Development aid: is the training forward / backward bitwise repeatable?"""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch, torch.nn.functional as F
import transmar_gan_b200 as T
from transmar_gan_b200 import dp, ops
from oracle import ngswin_oracle as O
m = T.NGswin(); m.load_state_dict(O.synthetic_state_dict("W1", seed=1), strict=True); m = m.cuda().train()
x = O.synthetic_input(2, 64, 96, seed=999).cuda(); t = O.synthetic_input(2, 64, 96, seed=1000).cuda()
ys = [m(x).detach().clone() for _ in range(8)]
print("train-mode forward max diff vs run 0:", [float((y - ys[0]).abs().max()) for y in ys])
with torch.no_grad():
    ye = [m(x).clone() for _ in range(8)]
print("inference forward max diff vs run 0:", [float((y - ye[0]).abs().max()) for y in ye])
# backward repeatability on ONE graph, fixed upstream gradient
y = m(x)
gup = torch.randn_like(y)
gs = []
for i in range(6):
    for p in m.parameters():
        p.grad = None
    y.backward(gup, retain_graph=True)
    gs.append(torch.cat([p.grad.flatten() for p in m.parameters()]).clone())
print("backward rel diff vs run 0:", [float((g - gs[0]).norm() / gs[0].norm()) for g in gs])
print("tc_error", ops.tc_error())
