"""Development aid: two default-mode forwards at batch 32 x 416^2 (for an ncu pass over the second one)."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import transmar_gan_b200 as T
from oracle import ngswin_oracle as O
m = T.NGswin(); m.load_state_dict(O.synthetic_state_dict("W1", seed=1)); m = m.cuda().eval()
m.calibrate_precision()
x = (torch.rand(int(sys.argv[1]) if len(sys.argv) > 1 else 32, 1, 416, 416, device="cuda") * 2 - 1)
torch.cuda.synchronize()
torch.cuda.profiler.start()
with torch.no_grad():
    y = m(x)
torch.cuda.synchronize()
torch.cuda.profiler.stop()
print("ok", float(y.abs().mean()))
