"""Development aid: marginal error of block 0's attention half at "tf32" / "tf32c" on W3 (probe slices and the
phantoms the weights were fitted to), everything else on the fp32 kernels."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import transmar_gan_b200 as T
from transmar_gan_b200.ngswin import _probe_slices
gold = torch.load("tests/golden/w3_trained.pt", weights_only=False)
m = T.NGswin(); m.load_state_dict(gold["state_dict"], strict=True); m = m.cuda().eval()
inputs = {"probe": _probe_slices(96, 4242).cuda(), "phantom": gold["phantom_in"].cuda()}
with torch.no_grad():
    for name, x in inputs.items():
        m._set_plan(["fp32"] * 41); ref = m._forward_planned(x)
        for u in (0, 2):
            for lv in ("tf32", "tf32c"):
                plan = ["fp32"] * 41; plan[u] = lv
                m._set_plan(plan)
                print(name, "unit", u, lv, f"{(m._forward_planned(x) - ref).abs().max().item():.2e}")
        for plan, label in ((["tf32"] * 41, "all T"), (["tf32c"] + ["tf32"] * 40, "C + 40 T"),
                            (["tf32c", "tf32"] * 6 + ["tf32"] * 29, "C on enc1 attention"),
                            (["fp32"] + ["tf32"] * 40, "f + 40 T")):
            m._set_plan(plan)
            print(name, label, f"{(m._forward_planned(x) - ref).abs().max().item():.2e}")
