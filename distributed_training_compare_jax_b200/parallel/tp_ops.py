"""Autograd-aware tensor-parallel region markers for the PyTorch (oracle / gloo) path — the explicit
counterparts of the sharding constraints the reference hands to GSPMD (model/CausalSelfAttention.py:28-31,49-50,
model/MLP.py:17-18,23-24):
    copy_to_tp      forward identity, backward all-reduce   (input of a column-parallel matmul, collective X5)
    reduce_from_tp  forward all-reduce, backward identity   (output of a row-parallel matmul, collectives X3/X4)
and the vocab-parallel softmax cross-entropy (collective X7).

Sequence-parallel mode (``axis.seq_parallel``, the torch counterpart of ``parallel/tp_sp.py``): the residual stream holds
only this rank's sequences, so the two markers become the halves of the all-reduce —
    copy_to_tp      forward all-gather of the rows,  backward reduce-scatter
    reduce_from_tp  forward reduce-scatter,          backward all-gather
and ``slice_rows`` (after the embedding) keeps the owned rows in the forward pass and gathers dL/dh[0] in the backward
pass, so that the replicated embedding tables receive the gradient of ALL rows on every rank."""
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


def _all_gather_rows(x, axis):
    parts = [torch.empty_like(x) for _ in range(axis.size)]
    dist.all_gather(parts, x.contiguous(), group=axis.group)
    return torch.cat(parts, dim=0)


def _reduce_scatter_rows(x, axis):
    x = x.contiguous().clone()
    dist.all_reduce(x, group=axis.group)          # gloo has no reduce_scatter: reduce everything, keep the owned rows
    per = x.shape[0] // axis.size
    return x[axis.rank * per:(axis.rank + 1) * per].clone()


class _GatherRows(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, axis):
        ctx.axis = axis
        return _all_gather_rows(x, axis)

    @staticmethod
    def backward(ctx, g):
        return _reduce_scatter_rows(g, ctx.axis), None


class _ScatterRows(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, axis):
        ctx.axis = axis
        return _reduce_scatter_rows(x, axis)

    @staticmethod
    def backward(ctx, g):
        return _all_gather_rows(g, ctx.axis), None


class _SliceRows(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, axis):
        ctx.axis = axis
        per = x.shape[0] // axis.size
        return x[axis.rank * per:(axis.rank + 1) * per].clone()

    @staticmethod
    def backward(ctx, g):
        return _all_gather_rows(g, ctx.axis), None


def _sp(axis) -> bool:
    return axis.size > 1 and getattr(axis, "seq_parallel", False)


def copy_to_tp(x, axis):
    if axis.size == 1:
        return x
    return _GatherRows.apply(x, axis) if _sp(axis) else _CopyToTP.apply(x, axis)


def reduce_from_tp(x, axis):
    if axis.size == 1:
        return x
    return _ScatterRows.apply(x, axis) if _sp(axis) else _ReduceFromTP.apply(x, axis)


def slice_rows(x, axis):
    """Sequence-parallel mode only: keep this rank's sequences (dim 0); identity otherwise."""
    return _SliceRows.apply(x, axis) if _sp(axis) else x


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
