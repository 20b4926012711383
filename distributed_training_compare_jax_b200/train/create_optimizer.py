"""Optimizer factory (component C10; reference: train/create_optimizer.py:8-12 =
``optax.chain(clip_by_global_norm(grad_clip), adamw(lr, weight_decay))``).

Semantics reproduced exactly: the norm is taken over the WHOLE gradient tree
(``where(norm < c, g, g/norm*c)``, no eps); AdamW uses b1 0.9, b2 0.999, eps 1e-8 outside the sqrt, bias
correction, and decoupled weight decay on EVERY leaf (biases, LayerNorm, embeddings included), constant LR.

Implementation: the parameters live in one flat buffer (``ParamStore``), so the optimizer is three launches —
sum-of-squares partials, clip-factor finalisation (which also advances the device-side step counter) and one
multi-tensor AdamW kernel that also refreshes the bf16 compute copy and zeroes the gradient.  Under tensor
parallelism only the sharded region's sum of squares is all-reduced (replicated leaves are counted once);
under data parallelism the gradient is summed over the DP axis first (``parallel/dp.py`` replaces that by the
fused reduce-scatter + sharded AdamW + all-gather path).
"""
from __future__ import annotations

from dataclasses import dataclass

import torch

from ..config.schema import OptimConfig
from ..ops import functional as Fn
from ..ops import native
from ..ops import reference as R


@dataclass
class OptimizerSpec:
    """What ``create_optimizer`` returns: the (clip, adamw) chain description; ``bind`` attaches it to an engine."""
    cfg: OptimConfig

    def bind(self, engine, mesh, clip_scope: str = "global", clip_scale: float = 1.0, grad_reducer=None):
        cls = NativeOptimizer if engine.backend == "native" else TorchOptimizer
        return cls(self.cfg, engine, mesh, clip_scope, clip_scale, grad_reducer)


def create_optimizer(cfg: OptimConfig) -> OptimizerSpec:
    return OptimizerSpec(cfg)


class _BoundOptimizer:
    def __init__(self, cfg, engine, mesh, clip_scope, clip_scale, grad_reducer):
        self.cfg, self.engine, self.mesh, self.store = cfg, engine, mesh, engine.store
        # clip_scope: "global" = norm over every stage (all-reduced over the PP axis) | "stage" = stage-local norm
        # (reference PP behaviour, quirk Q3).  clip_scale multiplies the threshold (quirk Q4: 1/S).
        self.clip_scope = clip_scope
        self.max_norm = cfg.grad_clip * clip_scale
        self.grad_reducer = grad_reducer  # callable() that makes flat_grad hold the DP-summed gradient


class TorchOptimizer(_BoundOptimizer):
    def __init__(self, *a):
        super().__init__(*a)
        self.count = 0
        self.last_norm = None

    @torch.no_grad()
    def step(self) -> None:
        st, mesh = self.store, self.mesh
        if self.grad_reducer is not None:
            self.grad_reducer()
        elif mesh.dp.size > 1:
            mesh.dp.all_reduce(st.flat_grad)
        g = st.flat_grad
        sq_rep = (g[:st.rep_numel] ** 2).sum()
        sq_shard = (g[st.rep_numel:] ** 2).sum()
        mesh.tp.all_reduce(sq_shard)
        sq = sq_rep + sq_shard
        if self.clip_scope == "global":
            mesh.pp.all_reduce(sq)
        norm = sq.sqrt()
        self.last_norm = norm
        st.step += 1
        self.count = int(st.step.item())  # survives checkpoint resume
        c = self.cfg
        g.mul_(R.clip_factor(norm, self.max_norm) if self.max_norm > 0 else 1.0)
        R.adamw_step(st.flat_master, st.flat_m, st.flat_v, g, self.count, c.lr, c.b1, c.b2, c.eps, c.weight_decay)
        g.zero_()


class NativeOptimizer(_BoundOptimizer):
    def __init__(self, *a):
        super().__init__(*a)
        st = self.store
        dev = st.device
        self.n_rep = native.sumsq_num_partials(max(st.rep_numel, 1))
        self.n_shard = native.sumsq_num_partials(max(st.shard_numel, 1))
        # partials layout: [rep partials | reduced shard scalar]; the last slot is all-reduced over TP (and PP)
        self.partials = torch.zeros(self.n_rep + 1, device=dev, dtype=torch.float32)
        self.shard_partials = torch.zeros(self.n_shard, device=dev, dtype=torch.float32)
        self.scalars = torch.zeros(2, device=dev, dtype=torch.float32)  # [norm, clip factor]

    def step(self) -> None:
        st, mesh, c = self.store, self.mesh, self.cfg
        if self.grad_reducer is not None:
            self.grad_reducer()
        elif mesh.dp.size > 1:
            mesh.dp.all_reduce(st.flat_grad)
        g = st.flat_grad
        if st.rep_numel > 0:
            Fn.sumsq_partials(g[:st.rep_numel], self.partials)
        slot = self.partials[self.n_rep:]
        Fn.sumsq_partials(g[st.rep_numel:], self.shard_partials)
        Fn.reduce_sum(self.shard_partials, 1.0, slot)
        mesh.tp.all_reduce(slot)
        if self.clip_scope == "global" and mesh.pp.size > 1:
            Fn.reduce_sum(self.partials, 1.0, slot)  # fold everything into the slot, then sum over stages
            mesh.pp.all_reduce(slot)
            Fn.clip_finalize(slot, 1, 1.0, self.max_norm, self.scalars, st.step)
        else:
            Fn.clip_finalize(self.partials, self.n_rep + 1, 1.0, self.max_norm, self.scalars, st.step)
        Fn.adamw(st.flat_master, st.flat_m, st.flat_v, g, st.flat_compute, self.scalars, st.step, c.lr, c.b1, c.b2,
                 c.eps, c.weight_decay, grad_scale=1.0, zero_grad=True)

    @property
    def last_norm(self):
        return self.scalars[0]
