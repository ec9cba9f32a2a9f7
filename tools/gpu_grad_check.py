"""Development aid: per-tensor gradient error of the tf32-forward / fp32-backward training path."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch, torch.nn.functional as F
import transmar_gan_b200 as T
from oracle import ngswin_oracle as O
torch.backends.cuda.matmul.allow_tf32 = False
torch.backends.cudnn.allow_tf32 = False
gold = torch.load("tests/golden/ngswin_outputs.pt", weights_only=False)
for mode in ("fp32", "tf32"):
    T.ops.set_precision(mode)
    m = T.NGswin(); m.load_state_dict(O.synthetic_state_dict("W1", seed=1), strict=True); m = m.cuda().train()
    x = O.synthetic_input(2, 64, 64, seed=999).cuda(); t = O.synthetic_input(2, 64, 64, seed=1000).cuda()
    y = m(x)
    for lossname, lossfn in (("l1", F.l1_loss), ("mse", F.mse_loss)):
        m.zero_grad()
        loss = lossfn(m(x), t); loss.backward()
        named = dict(m.named_parameters())
        if lossname == "l1":
            ref = gold[("W1", "grads")]
        else:
            sd = {k: (v.cuda().double().requires_grad_(True) if v.is_floating_point() and not k.endswith("attn_mask") else v.cuda())
                  for k, v in O.synthetic_state_dict("W1", seed=1).items()}
            F.mse_loss(O.ngswin_forward(x.double(), sd), t.double()).backward()
            ref = {k: v.grad.float().cpu() for k, v in sd.items() if isinstance(v, torch.Tensor) and v.requires_grad and v.grad is not None}
        errs = []
        for n, g in ref.items():
            if g.norm() < 1e-12: continue
            errs.append((((named[n].grad.cpu().double() - g.double()).norm() / g.double().norm()).item(), n, g.norm().item()))
        errs.sort(reverse=True)
        print(mode, lossname, "loss", loss.item(), "worst:", [(f"{e:.3e}", n, f"{gn:.2e}") for e, n, gn in errs[:5]], "median", errs[len(errs)//2][0], flush=True)
