"""MLP block (component C7; reference: model/MLP.py:12-26): fc1 → tanh-GELU → fc2.

Tensor parallel: fc1 is column-parallel (local width F/tp), fc2 row-parallel; the partial fc2 outputs are
summed across the TP axis (reference: the ``with_sharding_constraint(..., P(None,None,None))`` at MLP.py:23-24).
Two execution paths share the parameters:
  * ``__call__``            PyTorch fp32 ops with autograd (oracle / CPU / gloo path)
  * ``native_forward/backward``  hand-written sm_100a kernels: fc1 GEMM with bias+GELU epilogue (keeps the
    pre-activation for the backward), fc2 GEMM with bias+residual epilogue, dgrad with GELU' epilogue,
    split-K wgrads with fp32 reduce-add, column-sum bias gradients.
"""
from __future__ import annotations

import torch

from ..ops import functional as Fn
from ..ops import reference as R
from ..parallel import tp_ops


class MLP:
    def __init__(self, config, local, mesh):
        self.config, self.local, self.mesh = config, local, mesh

    # ------------------------------------------------------------------ torch path
    def __call__(self, p: dict, l: int, x: torch.Tensor) -> torch.Tensor:
        tp = self.mesh.tp
        x = tp_ops.copy_to_tp(x, tp)
        h = R.gelu_tanh(x @ p["w1"][l] + p["b1"][l])
        y = tp_ops.reduce_from_tp(h @ p["w2"][l], tp)
        return y + p["b2"][l]

    # ------------------------------------------------------------------ native path
    def native_forward(self, eng, ws, l: int) -> None:
        """ws.ln2[l] -> ws.h[l+1] = ws.hmid[l] + fc2(gelu(fc1(ln2)))."""
        c, tp = eng.c, self.mesh.tp
        Fn.gemm(ws.ln2[l], c["w1"][l], b_mn=True, bias=c["b1"][l], epi="bias_gelu", out=ws.f1[l], out2=ws.g[l])
        eng.row_parallel_out(ws.g[l], c["w2"][l], c["b2"][l], residual=ws.hmid[l], out=ws.h[l + 1])

    def native_backward(self, eng, ws, l: int, dh: torch.Tensor) -> None:
        """dh = dL/dh[l+1].  Accumulates fc1/fc2 grads; leaves dL/d(ln2 out) in ws.dln (TP-reduced)."""
        c, g = eng.c, eng.g
        def fc2_param_grads():
            Fn.colsum(dh, g["b2"][l])
            Fn.gemm(ws.g[l], dh, a_mn=True, b_mn=True, epi="atomic_f32", out=g["w2"][l], splits=eng.wgrad_splits)

        def fc1_param_grads():
            Fn.colsum(ws.df1, g["b1"][l])
            Fn.gemm(ws.ln2[l], ws.df1, a_mn=True, b_mn=True, epi="atomic_f32", out=g["w1"][l], splits=eng.wgrad_splits)

        eng.side_work(fc2_param_grads, layer=l)
        Fn.gemm(dh, c["w2"][l], epi="dgelu", aux=ws.f1[l], out=ws.df1)
        eng.side_work(fc1_param_grads, layer=l)
        ws.dln_cur = eng.tp_reduced_gemm(ws.df1, c["w1"][l], out=ws.dln)
