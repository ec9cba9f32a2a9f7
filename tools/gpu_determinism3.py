"""Development aid: repeatability of the stage-level backward pieces (PatchMerging, pixel-shuffle+concat)
and of the whole-model backward per top-level module."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import transmar_gan_b200 as T
from oracle import ngswin_oracle as O
ops = T.ops
def rel(a, b): return float((a - b).norm() / (b.norm() + 1e-30))
g = torch.Generator().manual_seed(3)
B, ph, pw = 2, 64, 96
x = torch.randn(B, ph * pw, 64, generator=g).cuda(); dy = torch.randn(B, ph * pw // 4, 64, generator=g).cuda()
lnw = torch.randn(256, generator=g).cuda(); lnb = torch.randn(256, generator=g).cuda(); rw = (torch.randn(64, 256, generator=g) * .05).cuda()
outs = []
for i in range(5):
    G = [torch.zeros_like(t) for t in (lnw, lnb, rw)]; dx = torch.empty_like(x)
    ops.patch_merge_bwd(x, dy, [lnw, lnb, rw], dx, G, ph, pw)
    outs.append((dx.clone(), torch.cat([t.flatten() for t in G])))
print("patch_merge_bwd dx", ["%.1e" % rel(o[0], outs[0][0]) for o in outs[1:]], "G", ["%.1e" % rel(o[1], outs[0][1]) for o in outs[1:]])
dcat = torch.randn(B, ph * pw, 84, generator=g).cuda()
ps = [ops.pixel_shuffle_cat_bwd(dcat, ph, pw) for _ in range(4)]
print("pixel_shuffle_cat_bwd", [["%.1e" % rel(a, b) for a, b in zip(p, ps[0])] for p in ps[1:]])
# whole model: gradient wrt intermediate tensors via hooks
m = T.NGswin(); m.load_state_dict(O.synthetic_state_dict("W1", seed=1), strict=True); m = m.cuda().train()
xin = O.synthetic_input(2, 64, 96, seed=999).cuda()
y = m(xin); gup = torch.randn_like(y)
names = [n for n, _ in m.named_parameters()]
runs = []
for i in range(5):
    for p in m.parameters(): p.grad = None
    y.backward(gup, retain_graph=True)
    runs.append({n: p.grad.clone() for n, p in m.named_parameters()})
def group(n):
    parts = n.split(".")
    return ".".join(parts[:3]) if parts[1] == "blocks" else ".".join(parts[:2])
groups = {}
for n in names: groups.setdefault(group(n), []).append(n)
for gname, ns in groups.items():
    d = max(rel(torch.cat([r[n].flatten() for n in ns]), torch.cat([runs[0][n].flatten() for n in ns])) for r in runs[1:])
    print(f"{gname:40s} {d:.1e}")
