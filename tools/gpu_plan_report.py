"""Development aid: the measured precision plan (per-unit marginal errors) of W1 / W2 / W3."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import transmar_gan_b200 as T
from oracle import ngswin_oracle as O
torch.backends.cuda.matmul.allow_tf32 = False
torch.backends.cudnn.allow_tf32 = False
gold = torch.load("tests/golden/w3_trained.pt", weights_only=False)
for kind in ("W1", "W3", "W2"):
    m = T.NGswin()
    m.load_state_dict(gold["state_dict"] if kind == "W3" else O.synthetic_state_dict(kind, seed=1), strict=True)
    m = m.cuda().eval()
    r = m.calibrate_precision(tol=0.0)          # tol 0 forces the marginal pass
    print(kind, "all-tc probe err", f"{r['err_all_tensor_core']:.2e}")
    print("  marginal:", " ".join(f"{e:.1e}" for e in r["marginal_err"]))
    r = m.calibrate_precision()
    if r.get("marginal_err_compensated"):
        print("  compensated marginal:", " ".join("-" if e is None else f"{e:.1e}" for e in r["marginal_err_compensated"]))
    print("  plan:", "".join({"tf32": "T", "tf32c": "C"}.get(p, "f") for p in r["plan"]), f"err {r['err_plan']:.2e}")
    if kind == "W3":
        x = gold["phantom_in"].cuda()
        with torch.no_grad():
            n = 41
            m._set_plan(["fp32"] * n); ref = m._forward_planned(x)
            print("  fp32 kernels vs reference on phantom:", f"{(ref.cpu() - gold['phantom_out']).abs().max().item():.2e}")
            for u in range(n):
                plan = ["fp32"] * n; plan[u] = "tf32"
                m._set_plan(plan)
                print(f"  unit {u}: {(m._forward_planned(x) - ref).abs().max().item():.1e}", end="")
            print()
            m._set_plan(["tf32"] * n)
            print("  all-tc on phantom:", f"{(m._forward_planned(x) - ref).abs().max().item():.2e}")
