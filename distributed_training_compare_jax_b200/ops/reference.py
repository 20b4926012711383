"""Plain PyTorch fp32 restatement of every native op (test oracle T0 of SURVEY.md §4 and the CPU path).

The math follows SURVEY.md §2.7: LayerNorm eps 1e-6 with E[x²]−E[x]² variance, tanh-GELU, softmax with an
additive 0/−1e9 causal mask, mean softmax-CE with integer labels, optax clip_by_global_norm + adamw.
Reference citations: model/CausalSelfAttention.py:34-42, model/MLP.py:14, model/TransformerBlock.py:16,22,
train/create_train_step.py:32-34, train/create_optimizer.py:9-12.
"""
from __future__ import annotations

import math

import torch
import torch.nn.functional as F


def gelu_tanh(x: torch.Tensor) -> torch.Tensor:
    return 0.5 * x * (1.0 + torch.tanh(math.sqrt(2.0 / math.pi) * (x + 0.044715 * x ** 3)))


def gelu_tanh_grad(x: torch.Tensor) -> torch.Tensor:
    k0, k1 = math.sqrt(2.0 / math.pi), 0.044715
    u = k0 * (x + k1 * x ** 3)
    t = torch.tanh(u)
    return 0.5 * (1 + t) + 0.5 * x * (1 - t * t) * k0 * (1 + 3 * k1 * x * x)


def gemm(a, b, *, a_mn=False, b_mn=False, epi="store", alpha=1.0, bias=None, aux=None, out=None):
    A = a.float().t() if a_mn else a.float()
    Bm = b.float() if b_mn else b.float().t()
    acc = (A @ Bm) * alpha
    if epi == "atomic_f32":
        return out.float() + acc if out is not None else acc
    if epi == "dgelu":
        return acc * gelu_tanh_grad(aux.float())
    if bias is not None:
        acc = acc + bias.float()
    if epi == "bias_residual":
        return acc + aux.float()
    if epi == "bias_gelu":
        return acc, gelu_tanh(acc)
    return acc


def layernorm_fwd(x, gamma, beta, eps=1e-6):
    xf = x.float()
    mean = xf.mean(-1)
    var = (xf * xf).mean(-1) - mean * mean
    rstd = torch.rsqrt(var.clamp_min(0) + eps)
    y = (xf - mean[:, None]) * rstd[:, None] * gamma.float() + beta.float()
    return y, mean, rstd


def layernorm_bwd(dy, x, mean, rstd, gamma, dres=None):
    dyf, xf = dy.float(), x.float()
    xhat = (xf - mean[:, None]) * rstd[:, None]
    wdy = dyf * gamma.float()
    c1 = wdy.mean(-1, keepdim=True)
    c2 = (wdy * xhat).mean(-1, keepdim=True)
    dx = rstd[:, None] * (wdy - c1 - xhat * c2)
    if dres is not None:
        dx = dx + dres.float()
    return dx, (dyf * xhat).sum(0), dyf.sum(0)


def causal_attention(q, k, v):
    """q,k,v: [B,H,T,hd] fp32 -> [B,H,T,hd]; dense softmax with the additive 0/-1e9 mask of the reference."""
    T, hd = q.shape[-2], q.shape[-1]
    s = torch.einsum("bhtd,bhsd->bhts", q, k) * (hd ** -0.5)
    mask = torch.where(torch.tril(torch.ones(T, T, device=q.device)) == 1, 0.0, -1e9)
    p = torch.softmax(s + mask, dim=-1)
    return torch.einsum("bhts,bhsd->bhtd", p, v)


def attention_from_qkv(qkv, B, H, T, hd):
    """qkv [B*T, 3*H*hd] -> out [B*T, H*hd] (fp32 math)."""
    Dl = H * hd
    q, k, v = (qkv[:, i * Dl:(i + 1) * Dl].float().reshape(B, T, H, hd).permute(0, 2, 1, 3) for i in range(3))
    o = causal_attention(q, k, v)
    return o.permute(0, 2, 1, 3).reshape(B * T, Dl)


def cross_entropy_rows(logits, labels):
    lf = logits.float()
    return torch.logsumexp(lf, -1) - lf.gather(-1, labels.long()[:, None])[:, 0]


def cross_entropy_grad(logits, labels, gscale):
    lf = logits.float()
    p = torch.softmax(lf, -1)
    p[torch.arange(lf.shape[0], device=lf.device), labels.long()] -= 1.0
    return p * gscale


def clip_factor(norm: torch.Tensor, max_norm: float) -> torch.Tensor:
    """optax.clip_by_global_norm: where(norm < c, 1, c / norm)."""
    return torch.where(norm < max_norm, torch.ones_like(norm), max_norm / norm)


def adamw_step(p, m, v, g, step: int, lr, b1=0.9, b2=0.999, eps=1e-8, wd=0.0):
    """optax.adamw update (eps outside the sqrt, bias-corrected, decoupled weight decay on every element)."""
    m.mul_(b1).add_(g, alpha=1 - b1)
    v.mul_(b2).addcmul_(g, g, value=1 - b2)
    mhat = m / (1 - b1 ** step)
    vhat = v / (1 - b2 ** step)
    p.sub_(lr * (mhat / (vhat.sqrt() + eps) + wd * p))
    return p
