"""Development aid: is a forward under a given per-unit plan bitwise repeatable, and which unit breaks it?"""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import transmar_gan_b200 as T
from oracle import ngswin_oracle as O
m = T.NGswin(); m.load_state_dict(O.synthetic_state_dict("W1", seed=1), strict=True); m = m.cuda().eval()
SHAPE = tuple(int(v) for v in sys.argv[1:4]) if len(sys.argv) > 3 else (2, 96, 160)
x = O.synthetic_input(*SHAPE, seed=999).cuda()
print("shape", SHAPE)
def run(plan):
    m._set_plan(plan)
    with torch.no_grad():
        return m._forward_planned(x).clone()
def rep(name, plan):
    a = run(plan); a_cpu = a.cpu(); junk = torch.full((64, 1024, 1024), float("nan"), device="cuda"); del junk
    b = run(plan); c = run(plan)
    print(f"{name:28s} diff12 {(a - b).abs().max().item():.3e} diff23 {(b - c).abs().max().item():.3e} nan {torch.isnan(a).any().item() or torch.isnan(b).any().item()}", flush=True)
n = 41
rep("all fp32", ["fp32"] * n)
rep("all tf32", ["tf32"] * n)
rep("alt A (even tf32), glue fp32", ["tf32" if i % 2 == 0 else "fp32" for i in range(40)] + ["fp32"])
rep("alt B (odd tf32), glue tf32", ["fp32" if i % 2 == 0 else "tf32" for i in range(40)] + ["tf32"])
rep("blocks fp32, glue tf32", ["fp32"] * 40 + ["tf32"])
rep("blocks tf32, glue fp32", ["tf32"] * 40 + ["fp32"])
A = ["tf32" if i % 2 == 0 else "fp32" for i in range(40)] + ["fp32"]
B = ["fp32" if i % 2 == 0 else "tf32" for i in range(40)] + ["tf32"]
ya = run(A).cpu(); yb = run(B).cpu(); yb2 = run(B).cpu(); ya2 = run(A).cpu()
print("test sequence: B vs B", (yb - yb2).abs().max().item(), "A vs A", (ya - ya2).abs().max().item())
for u in range(41):
    plan = list(B); plan[u] = "fp32" if plan[u] == "tf32" else "tf32"
    a = run(plan); b = run(plan)
    if not torch.equal(a, b):
        print("  B with unit", u, "flipped: diff", (a - b).abs().max().item())
