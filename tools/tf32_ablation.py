"""Development aid (CPU): which block GEMM's TF32 operand rounding dominates the whole-model error?
Monkeypatches the oracle with truncation-to-TF32 at selected GEMM inputs."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch, torch.nn.functional as F
from oracle import ngswin_oracle as O

def tf32(x):  # round-to-nearest-even to 10 mantissa bits
    i = x.contiguous().view(torch.int32)
    i = (i + 0xFFF + ((i >> 13) & 1)) & ~0x1FFF
    return i.view(torch.float32)

def split(x):  # hi + lo, both tf32
    hi = tf32(x); return hi, tf32(x - hi)

MODE = {}
BLOCK_MODE = {}        # attention prefix -> {gemm name: mode}, overrides MODE for that block's attention GEMMs
_CUR = [None]
def mm(name, a, b):
    """a @ b with the configured rounding: 0 exact, 1 tf32 both, 2 = 3-pass split (hi*hi+hi*lo+lo*hi), 3 = lo only on a, 4 = lo only on b"""
    m = BLOCK_MODE.get(_CUR[0], {}).get(name, MODE.get(name, 0))
    if m == 0: return a @ b
    if m == 1: return tf32(a) @ tf32(b)
    ah, al = split(a); bh, bl = split(b)
    if m == 2: return ah @ bh + ah @ bl + al @ bh
    if m == 3: return ah @ bh + al @ bh
    if m == 4: return ah @ bh + ah @ bl

def window_attention(x, p, prefix, num_heads, win, mask=None):
    if win != 8: return ORIG(x, p, prefix, num_heads, win, mask)
    B_, N, C = x.shape
    _CUR[0] = prefix
    qkv = mm("qkv", x, p[prefix + "qkv.weight"].t()) + p[prefix + "qkv.bias"]
    qkv = qkv.reshape(B_, N, 3, num_heads, -1).permute(2, 0, 3, 1, 4)
    q, k, v = qkv[0], qkv[1], qkv[2]
    qn = q / q.norm(dim=-1, keepdim=True).clamp_min(1e-12)
    kn = k / k.norm(dim=-1, keepdim=True).clamp_min(1e-12)
    scale = p[prefix + "logit_scale"].clamp(max=O.LOGIT_SCALE_MAX).exp()
    attn = mm("qk", qn * scale, kn.transpose(-2, -1))
    table = p[prefix + "relative_position_bias_table"]
    idx = O.relative_position_index(win, win)
    attn = attn + table[idx.reshape(-1)].reshape(N, N, num_heads).permute(2, 0, 1).unsqueeze(0)
    if mask is not None:
        nW = mask.shape[0]
        attn = (attn.reshape(B_ // nW, nW, num_heads, N, N) + mask[None, :, None]).reshape(B_, num_heads, N, N)
    attn = attn.softmax(dim=-1)
    out = mm("pv", attn, v).transpose(1, 2).reshape(B_, N, -1)
    y = mm("proj", out, p[prefix + "proj.weight"].t()) + p[prefix + "proj.bias"]
    _CUR[0] = None
    return y

ORIG = O.window_attention
O.window_attention = window_attention
_lin = F.linear
def linear(x, w, b=None):
    name = {(128, 64): "fc1", (64, 128): "fc2"}.get(tuple(w.shape)) if x.shape[-1] in (64, 128) and x.dim() == 3 else None
    if name and MODE.get(name, 0):
        y = mm(name, x, w.t())
        return y + b if b is not None else y
    return _lin(x, w, b)
O.F.linear = linear

if __name__ == "__main__":
    import sys as _sys
    kinds = _sys.argv[1:] or ["W1", "W2"]
    for kind in kinds:
        if kind == "W3":      # trained-like weights on the structured phantom they were fitted to
            gold = torch.load(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden",
                                           "w3_trained.pt"), weights_only=False)
            sd, x = gold["state_dict"], gold["phantom_in"][:1]
        else:
            sd = O.synthetic_state_dict(kind, seed=1)
            x = O.synthetic_input(1, 64, 64, seed=999)
        with torch.no_grad():
            MODE.clear(); ref = O.ngswin_forward(x, sd)
            for cfg in ({"qkv": 1}, {"qk": 1}, {"pv": 1}, {"proj": 1}, {"fc1": 1}, {"fc2": 1},
                        dict(qkv=1, qk=1, pv=1, proj=1, fc1=1, fc2=1),
                        dict(qkv=2, qk=2, pv=2, proj=2, fc1=2, fc2=2),
                        dict(qkv=2, qk=1, pv=1, proj=1, fc1=1, fc2=1),
                        dict(qkv=2, qk=2, pv=1, proj=1, fc1=1, fc2=1),
                        dict(qkv=1, qk=1, pv=1, proj=1, fc1=2, fc2=2),
                        dict(qkv=2, qk=2, pv=2, proj=2, fc1=1, fc2=1),
                        dict(qkv=3, qk=1, pv=1, proj=1, fc1=1, fc2=1),
                        dict(qkv=4, qk=1, pv=1, proj=1, fc1=1, fc2=1),
                        dict(qkv=3, qk=3, pv=1, proj=1, fc1=1, fc2=1)):
                MODE.clear(); MODE.update(cfg)
                y = O.ngswin_forward(x, sd)
                print(kind, cfg, f"max-abs {(y - ref).abs().max().item():.2e}", flush=True)
