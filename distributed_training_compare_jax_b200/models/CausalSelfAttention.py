"""Causal self-attention (component C6; reference: model/CausalSelfAttention.py:13-52).

q/k/v projections are separate leaves in the parameter tree (q_proj/k_proj/v_proj) but stored and executed as
ONE fused [D, 3·D/tp] GEMM.  Tensor parallel: heads are split across the TP axis (column-parallel q/k/v,
reference CausalSelfAttention.py:28-31), out_proj is row-parallel and its partial sums are reduced over the TP
axis (reference :49-50).  The dense (B,H,T,T) softmax of the reference is replaced by tcgen05 flash attention
on the native path; the torch path keeps the reference's dense formulation with the additive 0/−1e9 mask.
"""
from __future__ import annotations

import torch

from ..ops import functional as Fn
from ..ops import reference as R
from ..parallel import tp_ops


class CausalSelfAttention:
    def __init__(self, config, local, mesh):
        self.config, self.local, self.mesh = config, local, mesh

    # ------------------------------------------------------------------ torch path
    def __call__(self, p: dict, l: int, x: torch.Tensor, mask: torch.Tensor | None = None) -> torch.Tensor:
        Hl, hd, Dl = self.local.heads_local, self.config.head_dim, self.local.d_local
        tp = self.mesh.tp
        x = tp_ops.copy_to_tp(x, tp)      # sequence-parallel mode: all-gather of the rows, so read the shape afterwards
        B, T, D = x.shape
        qkv = x @ p["wqkv"][l] + p["bqkv"][l]
        q, k, v = (qkv[..., i * Dl:(i + 1) * Dl].reshape(B, T, Hl, hd) for i in range(3))
        w = torch.einsum("bthd,bshd->bhts", q, k) * (hd ** -0.5)
        if mask is not None:
            w = w + mask
        w = torch.softmax(w, dim=-1)
        out = torch.einsum("bhts,bshd->bthd", w, v).reshape(B, T, Dl)
        out = tp_ops.reduce_from_tp(out @ p["wo"][l], tp)
        return out + p["bo"][l]

    # ------------------------------------------------------------------ native path
    def native_forward(self, eng, ws, l: int) -> None:
        """ws.ln1[l] -> ws.hmid[l] = ws.h[l] + out_proj(attention(qkv(ln1)))."""
        c = eng.c
        Fn.gemm(ws.ln1[l], c["wqkv"][l], b_mn=True, bias=c["bqkv"][l], out=ws.qkv[l])
        Fn.flash_attn_fwd(ws.qkv[l], ws.B, self.local.heads_local, ws.T, self.config.head_dim, out=ws.attn[l],
                          lse2=ws.lse2[l])
        eng.row_parallel_out(ws.attn[l], c["wo"][l], c["bo"][l], residual=ws.h[l], out=ws.hmid[l])

    def native_backward(self, eng, ws, l: int, dh: torch.Tensor) -> None:
        """dh = dL/dhmid[l].  Accumulates attention grads; leaves dL/d(ln1 out) in ws.dln (TP-reduced)."""
        c, g = eng.c, eng.g
        def out_param_grads():
            Fn.colsum(dh, g["bo"][l])
            Fn.gemm(ws.attn[l], dh, a_mn=True, b_mn=True, epi="atomic_f32", out=g["wo"][l], splits=eng.wgrad_splits)

        def qkv_param_grads():
            Fn.colsum(ws.dqkv, g["bqkv"][l])
            Fn.gemm(ws.ln1[l], ws.dqkv, a_mn=True, b_mn=True, epi="atomic_f32", out=g["wqkv"][l], splits=eng.wgrad_splits)

        eng.side_work(out_param_grads, layer=l)
        Fn.gemm(dh, c["wo"][l], out=ws.dattn)
        Fn.flash_attn_bwd(ws.qkv[l], ws.attn[l], ws.dattn, ws.lse2[l], ws.B, self.local.heads_local, ws.T,
                          self.config.head_dim, dqkv=ws.dqkv, dq_accum=ws.dq_accum, delta=ws.delta, dq_clean=True)
        eng.side_work(qkv_param_grads, layer=l)
        ws.dln_cur = eng.tp_reduced_gemm(ws.dqkv, c["wqkv"][l], out=ws.dln)
