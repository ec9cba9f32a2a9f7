"""TEST INFRASTRUCTURE — writes tests/golden/*.pt from the UNMODIFIED reference.

Run in the build container only (needs /root/reference):  python oracle/make_golden.py
The reference ships no golden vectors, so these are outputs of the reference
itself, imported through oracle/ref_shim.py, on seeded synthetic weights (W1, W2 of
oracle/ngswin_oracle.py::synthetic_state_dict) and inputs.  Only inputs' seeds, outputs and
a few intermediate module outputs are stored (small); weights are regenerated from the seed.
"""
from __future__ import annotations

import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from oracle import ngswin_oracle as O  # noqa: E402
from oracle import ref_shim  # noqa: E402

OUT = os.path.join(ROOT, "tests", "golden")


def main():
    torch.manual_seed(0)
    torch.set_num_threads(max(1, os.cpu_count() or 1))
    R = ref_shim.load_reference()
    model = R["NGswin"]().eval()
    ref_sd = model.state_dict()
    os.makedirs(OUT, exist_ok=True)

    # 1. the state_dict contract (names, shapes, dtypes, order) + integer buffers
    contract = [(k, tuple(v.shape), str(v.dtype)) for k, v in ref_sd.items()]
    torch.save({
        "contract": contract,
        "rel_index_8": R["get_relative_position_index"](8, 8),
        "rel_index_2": R["get_relative_position_index"](2, 2),
        "n_params": sum(p.numel() for p in model.parameters()),
    }, os.path.join(OUT, "state_dict_contract.pt"))

    # 2. integer index maps of partition / roll / unpartition (window_utils.py:155-181,
    #    transformer_blocks.py:109) pushed through the reference functions with arange tensors
    maps = {}
    for (B, ph, pw) in ((1, 16, 16), (2, 24, 40), (1, 64, 64)):
        ids = torch.arange(B * ph * pw, dtype=torch.float64).reshape(B, ph, pw, 1)
        for s in (0, 4):
            rolled = torch.roll(ids, shifts=(-s, -s), dims=(1, 2)) if s else ids
            win, (wh, ww) = R["window_partition"](rolled, 8)
            back = R["window_unpartition"](win, (wh, ww))
            back = torch.roll(back, shifts=(s, s), dims=(1, 2)) if s else back
            assert torch.equal(back, ids)
            maps[(B, ph, pw, s)] = win.reshape(-1, 64).to(torch.int64)
    # shift masks as the reference's NSTB.make_mask builds them (transformer_blocks.py:63-81)
    blk = model.encoder_layer1.blocks[1]
    masks = {(ph, pw): blk.make_mask((ph, pw)).to(torch.int8) for (ph, pw) in
             ((16, 16), (24, 40), (8, 16), (64, 64))}
    torch.save({"partition_maps": maps, "masks_i8": masks}, os.path.join(OUT, "index_maps.pt"))

    # 3. whole-model outputs + per-module intermediates
    gold = {}
    for kind in ("W1", "W2"):
        sd = O.synthetic_state_dict(kind, seed=1)
        model.load_state_dict(sd, strict=True)
        for (B, H, W) in ((1, 64, 64), (2, 96, 160), (1, 100, 70), (1, 128, 128)):
            x = O.synthetic_input(B, H, W, seed=999)
            with torch.no_grad():
                y = model(x)
            gold[(kind, B, H, W)] = y.clone()
        # per-module: NSTB (shifted + unshifted, 6 heads) and pieces, at 24x40 tokens
        g = torch.Generator().manual_seed(5)
        z = torch.randn(1, 24 * 40, 64, generator=g)
        for bi in (0, 1):
            blk = model.encoder_layer1.blocks[bi]
            with torch.no_grad():
                _, out, _ = blk(z, (24, 40))
                win, _ = blk.ngram_window_partition(z.view(1, 24, 40, 64))
                ctx = blk.ngram_window_partition.ngram_context(z.view(1, 24, 40, 64))
            gold[(kind, "nstb6", bi)] = out.clone()
            gold[(kind, "nwp6", bi)] = win.clone()
            gold[(kind, "ctx6", bi)] = ctx.reshape(1, 3, 5, 64).clone()
        for bi in (0, 1):
            blk = model.encoder_layer2.blocks[bi]
            with torch.no_grad():
                _, out, _ = blk(z, (24, 40))
            gold[(kind, "nstb4", bi)] = out.clone()
        with torch.no_grad():
            pm, _ = model.encoder_layer1.downsample(z, (24, 40))
        gold[(kind, "patch_merge")] = pm.clone()
        gold[(kind, "z_seed")] = 5
        # gradients of the L1 loss wrt a few parameters and the input (config 3 recipe)
        model.zero_grad(set_to_none=True)
        x = O.synthetic_input(2, 64, 64, seed=999).requires_grad_(True)
        t = O.synthetic_input(2, 64, 64, seed=1000)
        loss = torch.nn.functional.l1_loss(model(x), t)
        loss.backward()
        # W2: all 504 parameter gradients; W1: blocks 0/1 of every stage + all non-block params
        grads = {n: p.grad.clone() for n, p in model.named_parameters()
                 if kind == "W2" or ".blocks." not in n or ".blocks.0." in n or ".blocks.1." in n}
        gold[(kind, "l1_loss")] = loss.detach().clone()
        gold[(kind, "grad_x")] = x.grad.clone()
        gold[(kind, "grads")] = grads
    gold[("pixel_shuffle", 2)] = R["pixel_shuffle_permute"](
        torch.arange(1 * 6 * 10 * 64, dtype=torch.float32).reshape(1, 60, 64), (6, 10), (12, 20))
    gold[("pixel_shuffle", 4)] = R["pixel_shuffle_permute"](
        torch.arange(1 * 3 * 5 * 64, dtype=torch.float32).reshape(1, 15, 64), (3, 5), (12, 20))
    torch.save(gold, os.path.join(OUT, "ngswin_outputs.pt"))
    sz = sum(os.path.getsize(os.path.join(OUT, f)) for f in os.listdir(OUT))
    print(f"wrote golden fixtures to {OUT} ({sz / 1e6:.2f} MB)")


if __name__ == "__main__":
    main()
