"""Pre-LN transformer block (component C5; reference: model/TransformerBlock.py:13-26):
x = x + Attn(LN0(x)); x = x + MLP(LN1(x)); LayerNorm eps 1e-6; no dropout inside the block.

On the native path the residual adds live in the out_proj / fc2 GEMM epilogues and the residual-gradient
adds in the LayerNorm backward kernel, so a block is 2 LN + 4 GEMM + 1 attention launches forward.
"""
from __future__ import annotations

import torch
import torch.nn.functional as F

from ..ops import functional as Fn
from .CausalSelfAttention import CausalSelfAttention
from .MLP import MLP


def layer_norm(x: torch.Tensor, scale: torch.Tensor, bias: torch.Tensor, eps: float) -> torch.Tensor:
    """flax.linen.LayerNorm: fast variance E[x^2]-E[x]^2."""
    mean = x.mean(-1, keepdim=True)
    var = (x * x).mean(-1, keepdim=True) - mean * mean
    return (x - mean) * torch.rsqrt(var.clamp_min(0) + eps) * scale + bias


class TransformerBlock:
    def __init__(self, config, local, mesh):
        self.config, self.local, self.mesh = config, local, mesh
        self.attn = CausalSelfAttention(config, local, mesh)
        self.mlp = MLP(config, local, mesh)

    # ------------------------------------------------------------------ torch path
    def __call__(self, p: dict, l: int, x: torch.Tensor, mask: torch.Tensor | None) -> torch.Tensor:
        eps = self.config.layernorm_eps
        x = x + self.attn(p, l, layer_norm(x, p["ln1_g"][l], p["ln1_b"][l], eps), mask)
        x = x + self.mlp(p, l, layer_norm(x, p["ln2_g"][l], p["ln2_b"][l], eps))
        return x

    # ------------------------------------------------------------------ native path
    def native_forward(self, eng, ws, l: int) -> None:
        c, eps = eng.c, self.config.layernorm_eps
        Fn.layernorm_fwd(ws.h[l], c["ln1_g"][l], c["ln1_b"][l], eps, out=ws.ln1[l], mean=ws.mean1[l], rstd=ws.rstd1[l])
        self.attn.native_forward(eng, ws, l)
        Fn.layernorm_fwd(ws.hmid[l], c["ln2_g"][l], c["ln2_b"][l], eps, out=ws.ln2[l], mean=ws.mean2[l], rstd=ws.rstd2[l])
        self.mlp.native_forward(eng, ws, l)

    def native_backward(self, eng, ws, l: int, dh_out: torch.Tensor, dh_mid: torch.Tensor, dh_in: torch.Tensor) -> None:
        """dh_out = dL/dh[l+1] (input); dh_mid is scratch; dh_in receives dL/dh[l]."""
        c, g = eng.c, eng.g
        self.mlp.native_backward(eng, ws, l, dh_out)
        Fn.layernorm_bwd(ws.dln_cur, ws.hmid[l], ws.mean2[l], ws.rstd2[l], c["ln2_g"][l], g["ln2_g"][l], g["ln2_b"][l],
                         dres=dh_out, out=dh_mid)
        self.attn.native_backward(eng, ws, l, dh_mid)
        Fn.layernorm_bwd(ws.dln_cur, ws.h[l], ws.mean1[l], ws.rstd1[l], c["ln1_g"][l], g["ln1_g"][l], g["ln1_b"][l],
                         dres=dh_mid, out=dh_in)
