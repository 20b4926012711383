"""Tensor-parallel fused GEMM + collective (components R3/R4; collectives X3/X4/X5/X8 of SURVEY.md §2.5 — the
all-reduces GSPMD derives from the reference's ``with_sharding_constraint(..., replicated)`` after the
row-parallel out_proj / fc2 matmuls, model/CausalSelfAttention.py:49-50 and model/MLP.py:23-24, and their
transposes in the backward pass).

Baseline (``fused_comm: false``): tcgen05 GEMM, then an NCCL all-reduce of the [B·T, D] result.
Fused (``fused_comm: true``): ONE kernel — the GEMM epilogue reduce-adds its accumulator tile straight into every
rank's copy of a symmetric bf16 buffer with ``multimem.red`` on the NVSwitch multicast address (in-switch
reduction + broadcast, tile by tile while later tiles are still in the tensor cores); rank 0's tiles also carry
bias + residual.  A flag barrier (tiny kernel, no host sync) then publishes the result.  Target buffers rotate
through a ring and are re-zeroed right after use, so a buffer is always clean — on every rank — one barrier
before anyone reduces into it again.
"""
from __future__ import annotations

import torch

from ..ops import functional as Fn
from .symm import FlagBarrier, SymmBuffer


class FusedTPHooks:
    RING = 2

    def __init__(self, engine, mesh):
        self.mesh = mesh
        tp = mesh.tp
        M, D = engine.micro_batch * engine.seq_len, engine.cfg.d_model
        self.M, self.D = M, D
        self.bufs = [SymmBuffer(M * D, torch.bfloat16, engine.device, tp.group) for _ in range(self.RING)]
        self.barrier = FlagBarrier(engine.device, tp.group)
        self.i = 0
        self.slot = 0
        if not all(b.mc_ptr for b in self.bufs):
            raise RuntimeError("fused TP needs NVSwitch multicast (NVLS) support for symmetric memory")

    def begin_step(self) -> None:
        """Start every step at ring position 0 behind a barrier: all buffers are clean on every rank, whatever the
        number of fused GEMMs per step (the sequence is baked into the captured CUDA graph)."""
        self.i = 0
        self.slot = 0
        self.barrier(FlagBarrier.N_SLOTS - 1)

    def _reduce_into(self, out: torch.Tensor, launch) -> None:
        buf = self.bufs[self.i % self.RING]
        self.i += 1
        launch((buf.mc_ptr, self.D))            # GEMM whose epilogue multimem.red-adds into all ranks' `buf`
        self.barrier(self.slot % (FlagBarrier.N_SLOTS - 1))   # all ranks' tiles have landed everywhere
        self.slot += 1
        out.copy_(buf.tensor.view(self.M, self.D))
        buf.tensor.zero_()                       # clean again long before the ring comes back to it

    def row_parallel_out(self, eng, x, w, bias, residual, out) -> None:
        lead = self.mesh.tp.rank == 0
        self._reduce_into(out, lambda mc: Fn.gemm(x, w, b_mn=True, epi="mc_red", bias=bias if lead else None,
                                                  aux=residual if lead else None, mc_out=mc))

    def reduced_gemm(self, eng, a, w, out) -> None:
        self._reduce_into(out, lambda mc: Fn.gemm(a, w, epi="mc_red", mc_out=mc))


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
