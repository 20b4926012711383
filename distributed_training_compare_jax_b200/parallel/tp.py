"""Tensor-parallel fused GEMM + collective (components R3/R4; collectives X3/X4/X5/X8 of SURVEY.md §2.5 — the
all-reduces GSPMD derives from the reference's ``with_sharding_constraint(..., replicated)`` after the
row-parallel out_proj / fc2 matmuls, model/CausalSelfAttention.py:49-50 and model/MLP.py:23-24, and their
transposes in the backward pass).

Baseline (``tp_fused: false``): tcgen05 GEMM, then an NCCL all-reduce of the [B·T, D] result.
Fused (``tp_fused: true``): ONE kernel — the GEMM epilogue reduce-adds its accumulator tile straight into every
rank's copy of the destination with ``multimem.red`` on the NVSwitch multicast address (in-switch reduction +
broadcast, tile by tile while later tiles are still in the tensor cores); rank 0's tiles also carry bias +
residual.  A flag barrier (tiny kernel, no host sync) then publishes the result.  Every TP-reduced activation of a
step (hmid[l], h[l+1], and the 2L+1 dX buffers of the backward pass) owns a slice of ONE symmetric arena that is
cleared with a single memset behind a barrier at the start of the step — so a fused op is GEMM + barrier, nothing
is copied and no buffer is recycled inside a step.
"""
from __future__ import annotations

import torch

from ..ops import functional as Fn
from .symm import FlagBarrier, SymmBuffer


class FusedTPHooks:
    def __init__(self, engine, mesh):
        self.mesh = mesh
        tp = mesh.tp
        L = engine.local.L
        M, D = engine.micro_batch * engine.seq_len, engine.cfg.d_model
        self.M, self.D = M, D
        n_buf = L + L + (2 * L + 1)            # hmid[l], h[l+1], dX of {fc1, qkv} per layer + lm_head
        self.arena = SymmBuffer(n_buf * M * D, torch.bfloat16, engine.device, tp.group)
        if not self.arena.mc_ptr:
            raise RuntimeError("fused TP needs NVSwitch multicast (NVLS) support for symmetric memory")
        self.barrier = FlagBarrier(engine.device, tp.group)
        views = self.arena.tensor.view(n_buf, M, D)
        ws = engine.ws[0]
        # the row-parallel outputs live in the arena from now on (the engine keeps using ws.hmid / ws.h as before)
        ws.hmid = views[:L]
        for l in range(L):
            ws.h[l + 1] = views[L + l]
        self.dx = [views[2 * L + i] for i in range(2 * L + 1)]
        self.mc = {}
        for i in range(n_buf):
            self.mc[views[i].data_ptr()] = self.arena.mc_ptr + i * M * D * 2
        self.dx_i = 0
        self.slot = 0

    def begin_step(self) -> None:
        """One memset of the whole arena, then a barrier: every destination of this step is clean on every rank."""
        self.arena.tensor.zero_()
        self.dx_i = 0
        self.slot = 0
        self.barrier(FlagBarrier.N_SLOTS - 1)

    def _publish(self) -> None:
        self.barrier(self.slot % (FlagBarrier.N_SLOTS - 1))   # all ranks' tiles have landed everywhere
        self.slot += 1

    def row_parallel_out(self, eng, x, w, bias, residual, out) -> None:
        lead = self.mesh.tp.rank == 0
        Fn.gemm(x, w, b_mn=True, epi="mc_red", bias=bias if lead else None, aux=residual if lead else None,
                mc_out=(self.mc[out.data_ptr()], self.D))
        self._publish()

    def reduced_gemm(self, eng, a, w, out) -> torch.Tensor:
        buf = self.dx[self.dx_i]
        self.dx_i += 1
        Fn.gemm(a, w, epi="mc_red", mc_out=(self.mc[buf.data_ptr()], self.D))
        self._publish()
        return buf


def attach_fused_tp(engine, mesh) -> None:
    """Enable the fused GEMM -> NVLS all-reduce path when the TP group has multicast-capable symmetric memory."""
    if engine.n_micro != 1:
        return
    try:
        engine.tp_hooks = FusedTPHooks(engine, mesh)
    except Exception as e:  # no NVLS / symmetric memory on this system: keep the NCCL baseline, loudly
        if mesh.is_main:
            print(f"[dtc] fused TP unavailable ({type(e).__name__}: {e}); using GEMM + NCCL all-reduce", flush=True)
        engine.tp_hooks = None
