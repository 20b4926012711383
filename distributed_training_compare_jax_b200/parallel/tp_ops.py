"""Autograd-aware tensor-parallel region markers for the PyTorch (oracle / gloo) path — the explicit
counterparts of the sharding constraints the reference hands to GSPMD (model/CausalSelfAttention.py:28-31,49-50,
model/MLP.py:17-18,23-24):
    copy_to_tp      forward identity, backward all-reduce   (input of a column-parallel matmul, collective X5)
    reduce_from_tp  forward all-reduce, backward identity   (output of a row-parallel matmul, collectives X3/X4)
and the vocab-parallel softmax cross-entropy (collective X7)."""
from __future__ import annotations

import torch
import torch.distributed as dist


class _CopyToTP(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, axis):
        ctx.axis = axis
        return x

    @staticmethod
    def backward(ctx, g):
        g = g.contiguous().clone()
        dist.all_reduce(g, group=ctx.axis.group)
        return g, None


class _ReduceFromTP(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, axis):
        x = x.contiguous().clone()
        dist.all_reduce(x, group=axis.group)
        return x

    @staticmethod
    def backward(ctx, g):
        return g, None


def copy_to_tp(x, axis):
    return x if axis.size == 1 else _CopyToTP.apply(x, axis)


def reduce_from_tp(x, axis):
    return x if axis.size == 1 else _ReduceFromTP.apply(x, axis)


class _VocabParallelCE(torch.autograd.Function):
    """Per-row loss from vocab-sharded logits: exchanges only (max, sum-exp, target-logit) per row."""

    @staticmethod
    def forward(ctx, logits, labels, vocab_start, axis):
        V = logits.shape[-1]
        m = logits.max(-1).values
        if axis.size > 1:
            dist.all_reduce(m, op=dist.ReduceOp.MAX, group=axis.group)
        e = torch.exp(logits - m[:, None])
        s = e.sum(-1)
        local = labels - vocab_start
        ok = (local >= 0) & (local < V)
        tgt = torch.where(ok, logits.gather(-1, local.clamp(0, V - 1)[:, None])[:, 0], torch.zeros_like(m))
        if axis.size > 1:
            dist.all_reduce(s, group=axis.group)
            dist.all_reduce(tgt, group=axis.group)
        ctx.save_for_backward(e / s[:, None], local, ok)
        return m + torch.log(s) - tgt

    @staticmethod
    def backward(ctx, g):
        p, local, ok = ctx.saved_tensors
        grad = p.clone()
        rows = torch.nonzero(ok)[:, 0]
        grad[rows, local[rows]] -= 1.0
        return grad * g[:, None], None, None, None


def vocab_parallel_cross_entropy(logits, labels, vocab_start, axis):
    return _VocabParallelCE.apply(logits, labels.long(), vocab_start, axis)
