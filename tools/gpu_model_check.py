"""Development aid: whole-model error of the tf32 / fp32 kernel modes against the CPU oracle."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import transmar_gan_b200 as T
from oracle import ngswin_oracle as O
torch.backends.cuda.matmul.allow_tf32 = False
torch.backends.cudnn.allow_tf32 = False
for kind in ("W1", "W2"):
    for seed in (1, 2):
        sd = O.synthetic_state_dict(kind, seed=seed)
        m = T.NGswin(); m.load_state_dict(sd, strict=True); m = m.cuda().eval()
        for (B, H, W) in ((1, 64, 64), (2, 128, 128), (1, 100, 70), (1, 416, 416)):
            if H == 416 and seed == 2: continue
            x = O.synthetic_input(B, H, W, seed=999)
            with torch.no_grad():
                yr = O.ngswin_forward(x, sd)
                res = []
                for mode in ("fp32", "tf32"):
                    T.ops.set_precision(mode)
                    y = m(x.cuda()).cpu()
                    res.append((y - yr).abs().max().item())
            print(f"{kind} seed {seed} {B}x{H}x{W}: max-abs vs oracle fp32-mode {res[0]:.2e} tf32-mode {res[1]:.2e} "
                  f"(|y| max {yr.abs().max().item():.3f}) tc_error {T.ops.tc_error()}", flush=True)
