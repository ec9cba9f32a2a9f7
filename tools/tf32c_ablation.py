"""Development aid (CPU): what does a COMPENSATED (hi + lo, 3-pass) qkv projection and q k^T in selected blocks buy
on the trained-like weights, every other GEMM of every block rounded to 11 bits?  (precision level "tf32c")"""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import torch
import tf32_ablation as A          # patches the oracle (its own __main__ loop runs for W1 by default: keep argv small)
from oracle import ngswin_oracle as O
gold = torch.load(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden",
                               "w3_trained.pt"), weights_only=False)
sd, x = gold["state_dict"], gold["phantom_in"]
allsix = dict(qkv=1, qk=1, pv=1, proj=1, fc1=1, fc2=1)
enc1 = [f"encoder_layer1.blocks.{i}.attn." for i in range(6)]
every = sorted({k[: k.index("attn.") + 5] for k in sd if ".attn.qkv.weight" in k})
with torch.no_grad():
    A.MODE.clear(); A.BLOCK_MODE.clear(); ref = O.ngswin_forward(x, sd)
    for name, blocks, comp in (("none", [], {}),
                               ("block 0: qkv+qk", enc1[:1], dict(qkv=2, qk=2)),
                               ("block 0: qkv only", enc1[:1], dict(qkv=2)),
                               ("block 0: all four exact", enc1[:1], dict(qkv=0, qk=0, pv=0, proj=0)),
                               ("blocks 0-1: qkv+qk", enc1[:2], dict(qkv=2, qk=2)),
                               ("enc1 0-5: qkv+qk", enc1, dict(qkv=2, qk=2)),
                               ("every block: qkv+qk", every, dict(qkv=2, qk=2))):
        A.MODE.clear(); A.MODE.update(allsix); A.BLOCK_MODE.clear()
        for b in blocks: A.BLOCK_MODE[b] = comp
        y = O.ngswin_forward(x, sd)
        print(f"{name:28s} max-abs {(y - ref).abs().max().item():.2e}  per slice",
              [f"{v:.1e}" for v in (y - ref).abs().amax((1, 2, 3)).tolist()], flush=True)
