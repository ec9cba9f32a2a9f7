"""The generator's training step as one replayable unit.

`L1TrainStep` is the baseline-variant recipe of the reference (`y = netG(x); F.l1_loss(y, t).backward();
optG.step()`, scripts/run_ablation_studies.py:1525-1535, Adam betas (.5, .999), lr 1e-4) on the drop-in
NGswin, with the data-parallel plumbing of `dp.py`:

    zero the flat gradient bucket -> forward -> L1 -> backward -> [all-reduce the bucket] -> Adam

A step at the reference's training crop (batch 16 x 128x128) is ~440 kernel launches, most of them tens of
microseconds long.  With `graph=True` the step is captured once into CUDA graphs (inputs are copied into
static buffers) and every later call is a replay: part A = zero + forward + loss + backward, part B = Adam.
On one process A and B are a single graph; with several ranks the NCCL all-reduce of the flat bucket runs
between the two graphs on the same stream.  Measured on a B200 the L1 step is GPU-bound either way (20.5 ms
replayed, 21.8 ms launched eagerly); the full GAN step gains more (33.2 vs 43.1 ms) because the stock
PyTorch discriminator adds several hundred small launches.
"""
from __future__ import annotations

from typing import Optional, Tuple

import torch
import torch.distributed as dist
import torch.nn.functional as F

from . import dp


def _adam(params, lr, betas, graph):
    """torch.optim.Adam as the reference configures it (training/train_combined.py:261-262); on CUDA
    parameters the fused multi-tensor implementation, capturable when the step is replayed from a graph."""
    params = list(params)
    on_cuda = bool(params) and all(p.is_cuda for p in params)
    return torch.optim.Adam(params, lr=lr, betas=betas, capturable=bool(graph), fused=True if on_cuda else None)


def _fused_acc():
    from . import ops
    return ops.fused_grad_accumulation(True)


def _calibrate(model):
    """Precision "auto": the per-unit plan is measured now - a CUDA-graph capture freezes whatever plan is in
    force (NGswin re-measures it every AUTO_RECALIBRATE_EVERY eager training forwards, never during replay)."""
    from . import ops
    if ops.get_precision() == "auto" and hasattr(model, "calibrate_precision") and next(model.parameters()).is_cuda:
        if getattr(model, "_auto_plan", None) is None or model._weights_key() != model._auto_key:
            model.calibrate_precision()


class L1TrainStep:
    """Callable `step(x, target) -> loss` (a 0-d CUDA tensor; reading it synchronises)."""

    def __init__(self, model: torch.nn.Module, lr: float = 1e-4, betas: Tuple[float, float] = (0.5, 0.999),
                 graph: bool = True, group=None, warmup: int = 3):
        self.model, self.group, self.use_graph = model, group, graph
        self.bucket = dp.FlatGradBucket(model.parameters())
        # capturable: the step counter lives on the device, so the update can be replayed from a graph
        self.opt = _adam(model.parameters(), lr, betas, graph)
        self.world = dist.get_world_size(group) if dist.is_available() and dist.is_initialized() else 1
        self._warmup = max(int(warmup), 1)
        self._x = self._t = self._loss = None
        self._gA: Optional[torch.cuda.CUDAGraph] = None
        self._gB: Optional[torch.cuda.CUDAGraph] = None

    # -- the recipe -----------------------------------------------------------------------
    def _fwd_bwd(self, x, t):
        self.bucket.zero()
        loss = F.l1_loss(self.model(x), t)
        with _fused_acc():                 # the kernels add their parameter gradients into the flat bucket
            loss.backward()
        return loss.detach()

    def _eager(self, x, t):
        loss = self._fwd_bwd(x, t)
        self.bucket.all_reduce(self.group)
        self.opt.step()
        return loss

    # -- capture --------------------------------------------------------------------------
    def _capture(self, x, t):
        if not x.is_cuda:
            raise RuntimeError("L1TrainStep: inputs must be CUDA tensors (there is no CPU path)")
        self._x, self._t = x.clone(), t.clone()
        _calibrate(self.model)
        # the warm-up steps must not count: parameters and optimizer state are restored in place afterwards
        params = [p for g in self.opt.param_groups for p in g["params"]]
        snap_p = [p.detach().clone() for p in params]
        snap_s = {p: {k: v.clone() for k, v in self.opt.state[p].items() if torch.is_tensor(v)}
                  for p in params if p in self.opt.state and self.opt.state[p]}
        side = torch.cuda.Stream()
        side.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(side):                 # warm-up: lazy kernel attributes, Adam state, cuDNN plans
            for _ in range(self._warmup):
                self._eager(self._x, self._t)
        torch.cuda.current_stream().wait_stream(side)
        with torch.no_grad():
            for p, q in zip(params, snap_p):
                p.copy_(q)
                for k, v in self.opt.state[p].items():
                    if torch.is_tensor(v):
                        v.copy_(snap_s[p][k]) if p in snap_s else v.zero_()
        torch.cuda.synchronize()
        self._gA = torch.cuda.CUDAGraph()
        if self.world == 1:
            with torch.cuda.graph(self._gA):
                self._loss = self._fwd_bwd(self._x, self._t)
                self.opt.step()
        else:
            with torch.cuda.graph(self._gA):
                self._loss = self._fwd_bwd(self._x, self._t)
            self._gB = torch.cuda.CUDAGraph()
            with torch.cuda.graph(self._gB, pool=self._gA.pool()):
                self.opt.step()

    def __call__(self, x: torch.Tensor, t: torch.Tensor) -> torch.Tensor:
        if not self.use_graph:
            return self._eager(x, t)
        if self._gA is None:
            self._capture(x, t)
        if x.shape != self._x.shape or t.shape != self._t.shape:
            raise RuntimeError(f"L1TrainStep was captured for input {tuple(self._x.shape)}; got {tuple(x.shape)} "
                               "(build another step object for another shape)")
        if x.data_ptr() != self._x.data_ptr():
            self._x.copy_(x, non_blocking=True)
        if t.data_ptr() != self._t.data_ptr():
            self._t.copy_(t, non_blocking=True)
        self._gA.replay()
        if self._gB is not None:
            self.bucket.all_reduce(self.group)
            self._gB.replay()
        return self._loss.clone()       # the static loss buffer is overwritten by the next replay


class GANTrainStep:
    """The full TransMAR-GAN step (training/train_combined.py:399-500; BASELINE config 4):

        D step: fake = netG(ct) without grad; hinge loss of D(cat[ct, real]) / D(cat[ct, fake]); Adam (lrD)
        G step: fake = netG(ct); logits, features = D(cat[ct, fake]); real features without grad;
                L = 0.1 hinge_g + 10 FM + 1 metal-aware L1 + 0.2 metal-aware edge + 0.5 metal consistency
                (+ 0.02 physics in the reference: torch-radon is an absent submodule there, so that term is
                omitted); Adam (lrG)

    = 2 generator forwards + 1 generator backward (this repo's kernels) and 4 discriminator forwards + 2
    backwards (stock PyTorch / cuDNN, `gan.py`).  Data-parallel: one flat gradient bucket per network, each
    all-reduced once per step.  During the G step the discriminator's parameters do not require grad (the
    reference accumulates and then discards those gradients, `netD.zero_grad()` :404); results are identical.
    With `graph=True` the step is replayed from CUDA graphs split at the two all-reduces.
    `step(ct, real) -> (loss_D, loss_G)` as 0-d CUDA tensors.

    `fused_losses` (default: on for CUDA modules): the metal-aware L1, edge and consistency terms and their
    gradient come from ONE kernel (`ops.MetalLossFunction`) instead of six stock
    pooling / elementwise passes and their backward; same arithmetic, summation order aside.
    `reuse_fake` (default off): the generator runs ONCE per step, with grad, and the
    D step sees `fake.detach()`.  The generator's weights do not change between the two forwards of the
    reference recipe and it has no dropout, so the result is the same function of the same weights; only the
    kernels differ (the no-grad forward runs the inference kernels), hence opt-in.
    `projector` (default None = the term is left out): a `radon.Radon(size, angles)`; the G loss gains
    `lambda_phys * physics_loss_syn(fake, real, M, projector)` as at train_combined.py:480,495."""

    LAMBDAS = dict(adv=0.1, fm=10.0, rec=1.0, edge=0.2, metal=0.5)        # train_combined.py:108-113
    METAL = dict(threshold=0.6, radius=5, beta=1.0, w_max=3.0)            # :116-119

    def __init__(self, netG: torch.nn.Module, netD: torch.nn.Module, lrG: float = 1e-4, lrD: float = 2e-4,
                 betas: Tuple[float, float] = (0.5, 0.999), graph: bool = True, group=None, warmup: int = 3,
                 fused_losses: Optional[bool] = None, reuse_fake: bool = False, projector=None,
                 lambda_phys: float = 0.02, split_graphs: Optional[bool] = None):
        from . import gan
        self.gan, self.netG, self.netD, self.group, self.use_graph = gan, netG, netD, group, graph
        on_cuda = next(netG.parameters()).is_cuda
        self.fused_losses = on_cuda if fused_losses is None else bool(fused_losses)
        self.reuse_fake = bool(reuse_fake)
        self.projector, self.lambda_phys = projector, float(lambda_phys)      # train_combined.py:113,269-272
        self._fake = None
        self.bucketG = dp.FlatGradBucket(netG.parameters())
        self.bucketD = dp.FlatGradBucket(netD.parameters())
        self.optG = _adam(netG.parameters(), lrG, betas, graph)
        self.optD = _adam(netD.parameters(), lrD, betas, graph)
        self.world = dist.get_world_size(group) if dist.is_available() and dist.is_initialized() else 1
        self._warmup = max(int(warmup), 1)
        self._graphs = None
        self._ct = self._real = self._lD = self._lG = None
        # graph mode with several ranks (or split_graphs=True, which the single-GPU test uses): the step is captured
        # in pieces around the two all-reduces, and the discriminator's all-reduce + Adam update run on a side stream
        # UNDER the generator forward of the G step, which does not depend on the discriminator
        self._split = (self.world > 1) if split_graphs is None else bool(split_graphs)
        self._side = None

    # -- the recipe, in the three pieces the all-reduces separate ---------------------------
    def _d_fwd_bwd(self, ct, real):
        self.bucketD.zero()
        if self.reuse_fake:
            self._fake = self.netG(ct)             # with grad: the G step below backpropagates through it
            fake = self._fake.detach()
        else:
            with torch.no_grad():
                fake = self.netG(ct)
        real_logits, _ = self.netD(torch.cat([ct, real], 1), return_features=False)
        fake_logits, _ = self.netD(torch.cat([ct, fake], 1), return_features=False)
        loss = self.gan.hinge_d_loss(real_logits, fake_logits)
        loss.backward()
        return loss.detach()

    def _d_step_g_fwd_bwd(self, ct, real):
        self.optD.step()
        return self._g_rest(ct, real, self._g_fwd(ct))

    def _g_fwd(self, ct):
        """Generator forward of the G step (independent of the discriminator and of its update)."""
        self.bucketG.zero()
        fake = self._fake if self.reuse_fake else self.netG(ct)
        self._fake = None
        return fake

    def _g_rest(self, ct, real, fake):
        """The rest of the G step: the (updated) discriminator on the fake, the losses, the generator backward."""
        gan, lam, mt = self.gan, self.LAMBDAS, self.METAL
        for p in self.bucketD.params:
            p.requires_grad_(False)
        try:
            fake_logits, fake_feats = self.netD(torch.cat([ct, fake], 1), return_features=True)
            with torch.no_grad():
                _, real_feats = self.netD(torch.cat([ct, real], 1), return_features=True)
            loss = (lam["adv"] * gan.hinge_g_loss(fake_logits)
                    + lam["fm"] * gan.feature_matching_loss(real_feats, fake_feats))
            if self.fused_losses:
                from . import ops
                image_terms, _ = ops.MetalLossFunction.apply(
                    fake, real, ct, lam["rec"], lam["edge"], lam["metal"], mt["threshold"], mt["beta"],
                    mt["radius"], mt["w_max"])
                loss = loss + image_terms
            else:
                M = gan.extract_metal_mask(ct, mt["threshold"])
                w = gan.compute_weight_map(ct, mt["beta"], mt["radius"], mt["w_max"], mt["threshold"])
                loss = (loss
                        + lam["rec"] * gan.compute_metal_aware_loss(fake, real, ct, mt["beta"], mt["radius"],
                                                                    mt["w_max"], mt["threshold"])
                        + lam["edge"] * gan.compute_metal_aware_edge_loss(fake, real, w)
                        + lam["metal"] * gan.metal_consistency_loss(fake, real, M))
            if self.projector is not None and self.lambda_phys != 0.0:
                M = gan.extract_metal_mask(ct, mt["threshold"])
                loss = loss + self.lambda_phys * gan.physics_loss_syn(fake, real, M, self.projector)
            with _fused_acc():
                loss.backward()
        finally:
            for p in self.bucketD.params:
                p.requires_grad_(True)
        return loss.detach()

    def _eager(self, ct, real):
        lD = self._d_fwd_bwd(ct, real)
        self.bucketD.all_reduce(self.group)
        lG = self._d_step_g_fwd_bwd(ct, real)
        self.bucketG.all_reduce(self.group)
        self.optG.step()
        return lD, lG

    # -- capture --------------------------------------------------------------------------
    def _snapshot(self):
        mods = (self.netG, self.netD)
        tensors = [t for m in mods for t in list(m.parameters()) + list(m.buffers())]
        return tensors, [t.detach().clone() for t in tensors]

    def _capture(self, ct, real):
        if not ct.is_cuda:
            raise RuntimeError("GANTrainStep: inputs must be CUDA tensors (there is no CPU path)")
        self._ct, self._real = ct.clone(), real.clone()
        _calibrate(self.netG)
        tensors, snap = self._snapshot()           # warm-up must not count (weights, spectral-norm u/v, Adam)
        # optimizer state that already exists (a resumed checkpoint) is restored; state the warm-up creates is zeroed
        snap_s = {id(p): {k: v.clone() for k, v in st.items() if torch.is_tensor(v)}
                  for opt in (self.optG, self.optD) for p, st in opt.state.items() if st}
        side = torch.cuda.Stream()
        side.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(side):
            for _ in range(self._warmup):
                self._eager(self._ct, self._real)
        torch.cuda.current_stream().wait_stream(side)
        with torch.no_grad():
            for t, s in zip(tensors, snap):
                t.copy_(s)
            for opt in (self.optG, self.optD):
                for p, st in opt.state.items():
                    old = snap_s.get(id(p))
                    for k, v in st.items():
                        if torch.is_tensor(v):
                            v.copy_(old[k]) if old is not None and k in old else v.zero_()
        torch.cuda.synchronize()
        g1 = torch.cuda.CUDAGraph()
        if not self._split:
            with torch.cuda.graph(g1):
                self._lD = self._d_fwd_bwd(self._ct, self._real)
                self._lG = self._d_step_g_fwd_bwd(self._ct, self._real)
                self.optG.step()
            self._graphs = (g1,)
        else:
            with torch.cuda.graph(g1):
                self._lD = self._d_fwd_bwd(self._ct, self._real)
            g2a = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g2a, pool=g1.pool()):
                fake = self._g_fwd(self._ct)
            gD = torch.cuda.CUDAGraph()             # replayed on the side stream CONCURRENTLY with g2a: its own memory
            with torch.cuda.graph(gD):              # pool, so that none of its temporaries aliases the forward's
                self.optD.step()
            g2b = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g2b, pool=g1.pool()):
                self._lG = self._g_rest(self._ct, self._real, fake)
            g3 = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g3, pool=g1.pool()):
                self.optG.step()
            self._side = torch.cuda.Stream()
            self._graphs = (g1, g2a, gD, g2b, g3)

    def __call__(self, ct: torch.Tensor, real: torch.Tensor):
        if not self.use_graph:
            return self._eager(ct, real)
        if self._graphs is None:
            self._capture(ct, real)
        if ct.shape != self._ct.shape or real.shape != self._real.shape:
            raise RuntimeError(f"GANTrainStep was captured for input {tuple(self._ct.shape)}; got {tuple(ct.shape)}")
        if ct.data_ptr() != self._ct.data_ptr():
            self._ct.copy_(ct, non_blocking=True)
        if real.data_ptr() != self._real.data_ptr():
            self._real.copy_(real, non_blocking=True)
        if len(self._graphs) == 1:
            self._graphs[0].replay()
        else:
            g1, g2a, gD, g2b, g3 = self._graphs
            cur = torch.cuda.current_stream()
            g1.replay()
            self._side.wait_stream(cur)
            with torch.cuda.stream(self._side):     # discriminator: all-reduce + Adam ...
                self.bucketD.all_reduce(self.group)
                gD.replay()
            g2a.replay()                            # ... under the generator forward
            cur.wait_stream(self._side)
            g2b.replay()
            self.bucketG.all_reduce(self.group)
            g3.replay()
        return self._lD.clone(), self._lG.clone()   # the static buffers are overwritten by the next replay
