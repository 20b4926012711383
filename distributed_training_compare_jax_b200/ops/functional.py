"""Python front end of the native sm_100a ops (thin, allocation-explicit wrappers around ``native.call``).

Every function takes torch CUDA tensors (bf16 activations / compute weights, fp32 statistics and gradients),
optionally writes into caller-provided outputs (the step engine passes preallocated buffers so a CUDA graph
replay never allocates) and enqueues exactly the kernels named in ``include/dtc_b200.h``.
"""
from __future__ import annotations

import threading

import torch

from . import native

EPI = {"store": 0, "bias_gelu": 1, "bias_residual": 2, "dgelu": 3, "atomic_f32": 4, "store_f32": 5, "store_stats": 8}
BF16 = torch.bfloat16
F32 = torch.float32


class gemm_impl:
    """Context manager selecting the GEMM kernel variant for the calls issued inside it by THIS host thread (the variant
    travels with every call — ``dtc_gemm_bf16_ex`` — so two threads or two engines with different policies do not
    interfere): 0 default policy, 1 one-tile-per-CTA, 2 / 3 CTA-pair kernels, 4 the 1-CTA persistent kernel."""
    _tls = threading.local()

    def __init__(self, impl: int):
        self.impl, self.prev = impl, 0

    @staticmethod
    def get() -> int:
        return getattr(gemm_impl._tls, "current", 0)

    def __enter__(self):
        self.prev = gemm_impl.get()
        gemm_impl._tls.current = self.impl

    def __exit__(self, *exc):
        gemm_impl._tls.current = self.prev


def _check_rowmajor(t: torch.Tensor, name: str) -> None:
    if t.dim() != 2 or t.stride(1) != 1:
        raise ValueError(f"{name}: expected a 2-D tensor with unit inner stride, got shape {tuple(t.shape)} "
                         f"stride {t.stride()}")


def gemm(a: torch.Tensor, b: torch.Tensor, *, a_mn: bool = False, b_mn: bool = False, epi: str = "store",
         alpha: float = 1.0, bias: torch.Tensor | None = None, aux: torch.Tensor | None = None,
         out: torch.Tensor | None = None, out2: torch.Tensor | None = None, splits: int = 0, stages: int = 0,
         block_n: int = 0, stats: torch.Tensor | None = None, n_valid: int = 0) -> torch.Tensor:
    """D[M,N] = A·Bᵀ on the tcgen05 GEMM.  ``a`` is [M,K] (or [K,M] when ``a_mn``), ``b`` is [N,K]
    (or [K,N] when ``b_mn``).  See ``dtc_gemm_bf16`` for the epilogues.  ``epi="store_stats"``: besides the bf16 output,
    ``stats`` (fp32 [M, 2·ceil(N / block_n), 2]) receives {max, Σ exp(x − max)} of every 128- (block_n 256) or 64-column
    (block_n 128) part of each row over the columns < ``n_valid`` (see :func:`ce_bwd_from_parts`)."""
    _check_rowmajor(a, "a")
    _check_rowmajor(b, "b")
    K, M = (a.shape[0], a.shape[1]) if a_mn else (a.shape[1], a.shape[0])
    Kb, N = (b.shape[0], b.shape[1]) if b_mn else (b.shape[1], b.shape[0])
    if K != Kb:
        raise ValueError(f"gemm: reduction dims differ ({K} vs {Kb})")
    mode = EPI[epi]
    if out is None:
        if mode == 4:
            raise ValueError("gemm: the fp32 reduce-add epilogue needs a caller-provided accumulator")
        out = torch.empty((M, N), device=a.device, dtype=F32 if mode == 5 else BF16)
    _check_rowmajor(out, "out")
    if mode == 1 and out2 is None:
        out2 = torch.empty((M, N), device=a.device, dtype=BF16)
    if mode == 8:
        block_n = block_n or 256
        n_parts = 2 * (-(-N // block_n))
        if stats is None or stats.dtype != F32 or stats.numel() < M * n_parts * 2 or not stats.is_contiguous():
            raise ValueError(f"gemm: store_stats needs a contiguous fp32 stats tensor of at least [{M}, {n_parts}, 2]")
        native.call("dtc_gemm_bf16_ex", a, a.stride(0), int(a_mn), b, b.stride(0), int(b_mn), M, N, K, mode, float(alpha),
                    bias, None, 0, out, out.stride(0), stats, int(n_valid or N), 0, stages, block_n, gemm_impl.get())
        return out
    native.call("dtc_gemm_bf16_ex", a, a.stride(0), int(a_mn), b, b.stride(0), int(b_mn), M, N, K, mode, float(alpha),
                bias, aux, aux.stride(0) if aux is not None else 0, out, out.stride(0), out2,
                out2.stride(0) if out2 is not None else 0, splits, stages, block_n, gemm_impl.get())
    return out if mode != 1 else (out, out2)


def gemm_scatter(a: torch.Tensor, b: torch.Tensor, peer_ptrs, n_ranks: int, rank: int, ldo: int, *, b_mn: bool = False,
                 stages: int = 0, block_n: int = 0) -> None:
    """GEMM -> reduce-scatter (sequence-parallel TP): D = A·Bᵀ whose row block of rank ``o`` is stored into slot ``rank``
    of rank ``o``'s partial-sum buffer.  ``peer_ptrs``: ctypes array of the n_ranks buffer addresses (each [M, ldo] bf16)."""
    _check_rowmajor(a, "a")
    _check_rowmajor(b, "b")
    M, K = a.shape
    Kb, N = (b.shape[0], b.shape[1]) if b_mn else (b.shape[1], b.shape[0])
    if K != Kb:
        raise ValueError(f"gemm_scatter: reduction dims differ ({K} vs {Kb})")
    native.call("dtc_gemm_bf16_scatter", a, a.stride(0), b, b.stride(0), int(b_mn), M, N, K, peer_ptrs, n_ranks, rank, ldo,
                stages, block_n)


def gemm_scatter_signal(a: torch.Tensor, b: torch.Tensor, peer_ptrs, n_ranks: int, rank: int, ldo: int, sig_ptrs, counters,
                        *, b_mn: bool = False, stages: int = 0, block_n: int = 0) -> None:
    """:func:`gemm_scatter` whose last CTA also publishes the tiles to every rank (folded barrier, experimental)."""
    _check_rowmajor(a, "a")
    _check_rowmajor(b, "b")
    M, K = a.shape
    Kb, N = (b.shape[0], b.shape[1]) if b_mn else (b.shape[1], b.shape[0])
    if K != Kb:
        raise ValueError(f"gemm_scatter: reduction dims differ ({K} vs {Kb})")
    native.call("dtc_gemm_bf16_scatter_signal", a, a.stride(0), b, b.stride(0), int(b_mn), M, N, K, peer_ptrs, n_ranks, rank,
                ldo, stages, block_n, sig_ptrs, counters)


def sp_ln_fwd_sync(res, bias, part, n_part: int, gamma, beta, hsum, mean, rstd, mc_out: int, peer_ptrs, n_out: int,
                   out_row0: int, eps: float, sig_ptrs, counters, rank: int, world: int, wait_in: bool,
                   signal_out: bool) -> None:
    """:func:`sp_ln_fwd` with the cross-rank waits / signals inside the kernel (folded barrier, experimental)."""
    Ms, D = res.shape
    native.call("dtc_sp_ln_fwd_sync", res, bias, part, n_part, gamma, beta, hsum, mean, rstd, mc_out, peer_ptrs, n_out,
                out_row0, Ms, D, float(eps), sig_ptrs, counters, rank, world, int(wait_in), int(signal_out))


def sp_ln_bwd_sync(part, n_part: int, x, mean, rstd, gamma, dres, dgamma, dbeta, mc_out: int, peer_ptrs, n_out: int,
                   out_row0: int, sig_ptrs, counters, rank: int, world: int, wait_in: bool, signal_out: bool) -> None:
    Ms, D = x.shape
    native.call("dtc_sp_ln_bwd_sync", part, n_part, x, mean, rstd, gamma, dres, dgamma, dbeta, mc_out, peer_ptrs, n_out,
                out_row0, Ms, D, sig_ptrs, counters, rank, world, int(wait_in), int(signal_out))


def sp_wait(sig_ptrs, counters, rank: int, world: int) -> None:
    """One tiny CTA that waits for the next signal of every rank (in front of a stock GEMM reading a gathered operand)."""
    native.call("dtc_sp_wait", sig_ptrs, counters, rank, world)


def sp_ln_fwd(res, bias, part, n_part: int, gamma, beta, hsum, mean, rstd, mc_out: int, peer_ptrs, n_out: int,
              out_row0: int, eps: float = 1e-6) -> None:
    """Owned rows: hsum = res + bias + Σ part[p]; LayerNorm(hsum) pushed to rows [out_row0, out_row0+Ms) of every rank's
    gathered buffer (``mc_out`` multicast address, or the ``peer_ptrs`` list when 0).  See ``dtc_sp_ln_fwd``."""
    Ms, D = res.shape
    native.call("dtc_sp_ln_fwd", res, bias, part, n_part, gamma, beta, hsum, mean, rstd, mc_out, peer_ptrs, n_out,
                out_row0, Ms, D, float(eps))


def sp_ln_bwd(part, n_part: int, x, mean, rstd, gamma, dres, dgamma, dbeta, mc_out: int, peer_ptrs, n_out: int,
              out_row0: int) -> None:
    """Owned rows: dx = LayerNorm'(Σ part[p]) + dres, pushed to every rank's gathered gradient buffer; dgamma / dbeta
    accumulate over the owned rows.  See ``dtc_sp_ln_bwd``."""
    Ms, D = x.shape
    native.call("dtc_sp_ln_bwd", part, n_part, x, mean, rstd, gamma, dres, dgamma, dbeta, mc_out, peer_ptrs, n_out,
                out_row0, Ms, D)


def layernorm_fwd(x, gamma, beta, eps: float = 1e-6, out=None, mean=None, rstd=None):
    M, D = x.shape
    out = torch.empty_like(x) if out is None else out
    mean = torch.empty(M, device=x.device, dtype=F32) if mean is None else mean
    rstd = torch.empty(M, device=x.device, dtype=F32) if rstd is None else rstd
    native.call("dtc_layernorm_fwd", x, gamma, beta, out, mean, rstd, M, D, float(eps))
    return out, mean, rstd


def layernorm_bwd(dy, x, mean, rstd, gamma, dgamma, dbeta, dres=None, out=None):
    M, D = x.shape
    out = torch.empty_like(x) if out is None else out
    native.call("dtc_layernorm_bwd", dy, x, mean, rstd, gamma, dres, out, dgamma, dbeta, M, D)
    return out


def embed_fwd(ids, wte, wpe, T: int, dropout_p: float, seed: int, step_ptr, stream_id: int = 0, row_offset: int = 0,
              out=None):
    M = ids.numel()
    V, D = wte.shape
    out = torch.empty((M, D), device=wte.device, dtype=BF16) if out is None else out
    native.call("dtc_embed_fwd", ids, wte, wpe, out, M, T, D, V, float(dropout_p), int(seed), step_ptr, int(stream_id),
                int(row_offset))
    return out


def embed_bwd(ids, dh, dwte, dwpe, T: int, dropout_p: float, seed: int, step_ptr, stream_id: int = 0,
              row_offset: int = 0):
    M = ids.numel()
    V, D = dwte.shape
    native.call("dtc_embed_bwd", ids, dh, dwte, dwpe, M, T, D, V, float(dropout_p), int(seed), step_ptr, int(stream_id),
                int(row_offset))


def ce_fwd_bwd(logits, labels, V: int, gscale: float, loss_rows=None):
    """In place: ``logits`` [M, ld] (bf16) become dlogits; returns the per-row losses (fp32 [M])."""
    M = logits.shape[0]
    loss_rows = torch.empty(M, device=logits.device, dtype=F32) if loss_rows is None else loss_rows
    native.call("dtc_ce_fwd_bwd", logits, logits.stride(0), labels, loss_rows, M, V, float(gscale))
    return loss_rows


def ce_bwd_from_parts(logits, labels, parts, n_parts: int, V: int, gscale: float, loss_rows=None):
    """Like :func:`ce_fwd_bwd`, with the row statistics taken from ``parts`` (written by ``gemm(epi="store_stats")``): one
    pass over the logits instead of two."""
    M = logits.shape[0]
    loss_rows = torch.empty(M, device=logits.device, dtype=F32) if loss_rows is None else loss_rows
    native.call("dtc_ce_bwd_from_parts", logits, logits.stride(0), labels, parts, n_parts, loss_rows, M, V, float(gscale))
    return loss_rows


def ce_stats(logits, labels, V_local: int, vocab_start: int, stats):
    native.call("dtc_ce_stats", logits, logits.stride(0), labels, stats, logits.shape[0], V_local, vocab_start)
    return stats


def ce_combine(stats, n_shards: int, M: int, lse, loss_rows):
    native.call("dtc_ce_combine", stats, n_shards, M, lse, loss_rows)


def ce_grad(logits, labels, lse, V_local: int, vocab_start: int, gscale: float):
    native.call("dtc_ce_grad", logits, logits.stride(0), labels, lse, logits.shape[0], V_local, vocab_start,
                float(gscale))


def reduce_sum(x, scale: float, out, accumulate: bool = False):
    native.call("dtc_reduce_sum", x, x.numel(), float(scale), out, int(accumulate))
    return out


def flash_attn_fwd(qkv: torch.Tensor, B: int, H: int, T: int, hd: int, out=None, lse2=None):
    """``qkv`` [B*T, 3*H*hd] bf16 with q | k | v column slices.  Returns (out [B*T, H*hd], lse2 [B,H,T])."""
    M = B * T
    Dl = H * hd
    out = torch.empty((M, Dl), device=qkv.device, dtype=BF16) if out is None else out
    lse2 = torch.empty((B, H, T), device=qkv.device, dtype=F32) if lse2 is None else lse2
    q, k, v = qkv[:, :Dl], qkv[:, Dl:2 * Dl], qkv[:, 2 * Dl:]
    native.call("dtc_flash_attn_fwd", q, k, v, qkv.stride(0), out, out.stride(0), lse2, B, H, T, hd)
    return out, lse2


def flash_attn_bwd(qkv, o, d_o, lse2, B: int, H: int, T: int, hd: int, dqkv=None, dq_accum=None, delta=None,
                   dq_clean: bool = False):
    """Returns dqkv [B*T, 3*H*hd] bf16.  ``dq_accum`` (fp32 [B*T, H*hd]) must be zero when the kernel starts: it is zeroed
    here, unless ``dq_clean`` — then the caller guarantees a zero buffer on entry and gets a zero buffer back (the final
    cast clears it behind its read: no fill kernel per layer)."""
    M, Dl = B * T, H * hd
    dqkv = torch.empty((M, 3 * Dl), device=qkv.device, dtype=BF16) if dqkv is None else dqkv
    if dq_accum is None:
        dq_accum = torch.zeros((M, Dl), device=qkv.device, dtype=F32)
    elif not dq_clean:
        dq_accum.zero_()
    delta = torch.empty((B, H, T), device=qkv.device, dtype=F32) if delta is None else delta
    q, k, v = qkv[:, :Dl], qkv[:, Dl:2 * Dl], qkv[:, 2 * Dl:]
    dk, dv = dqkv[:, Dl:2 * Dl], dqkv[:, 2 * Dl:]
    native.call("dtc_flash_attn_bwd", q, k, v, qkv.stride(0), o, o.stride(0), d_o, d_o.stride(0), lse2, delta, dq_accum,
                dq_accum.stride(0), dk, dv, dqkv.stride(0), B, H, T, hd)
    native.call("dtc_cast_f32_to_bf16_2d_clear" if dq_clean else "dtc_cast_f32_to_bf16_2d", dq_accum, dq_accum.stride(0),
                dqkv, dqkv.stride(0), M, Dl)
    return dqkv


def colsum(x, out):
    native.call("dtc_colsum", x, x.stride(0), x.shape[0], x.shape[1], out)
    return out


class GroupArgs:
    """Host-side argument block of a grouped launch (ctypes arrays built once, e.g. per layer at engine construction)."""

    def __init__(self, **arrays):
        import ctypes as C
        kinds = {"ptr": C.c_ulonglong, "ll": C.c_longlong, "int": C.c_int}
        self.count = None
        for name, (kind, values) in arrays.items():
            values = list(values)
            if self.count is None:
                self.count = len(values)
            elif self.count != len(values):
                raise ValueError("grouped launch: all argument lists must have the same length")
            setattr(self, name, (kinds[kind] * len(values))(*values))


def wgrad_group_args(xs, dys, outs) -> GroupArgs:
    """Argument block for :func:`wgrad_grouped`: out[p] [M_p, N_p] fp32 += xs[p]ᵀ · dys[p] with xs[p] [K, M_p], dys[p] [K, N_p]."""
    for x, dy, o in zip(xs, dys, outs):
        _check_rowmajor(x, "x"), _check_rowmajor(dy, "dy"), _check_rowmajor(o, "out")
        if x.shape[0] != dy.shape[0] or tuple(o.shape) != (x.shape[1], dy.shape[1]) or o.dtype != F32:
            raise ValueError(f"wgrad_grouped: shapes {tuple(x.shape)} {tuple(dy.shape)} -> {tuple(o.shape)} {o.dtype}")
    if len({x.shape[0] for x in xs}) != 1:
        raise ValueError("wgrad_grouped: all problems must share the reduction length")
    args = GroupArgs(a=("ptr", [x.data_ptr() for x in xs]), lda=("ll", [x.stride(0) for x in xs]),
                     b=("ptr", [d.data_ptr() for d in dys]), ldb=("ll", [d.stride(0) for d in dys]),
                     m=("int", [x.shape[1] for x in xs]), n=("int", [d.shape[1] for d in dys]),
                     out=("ptr", [o.data_ptr() for o in outs]), ldo=("ll", [o.stride(0) for o in outs]))
    args.K = xs[0].shape[0]
    return args


def wgrad_grouped(args: GroupArgs, alpha: float = 1.0, splits: int = 1, stages: int = 0, pair: bool = False) -> None:
    """The weight gradients of one layer (<= 4 GEMMs sharing K = tokens) as ONE persistent tcgen05 launch.
    ``pair``: 256 x 256 tiles on CTA pairs (cta_group::2) instead of 128 x 128 tiles on single CTAs."""
    native.call("dtc_gemm_wgrad_grouped_pair" if pair else "dtc_gemm_wgrad_grouped", args.count, args.a, args.lda, args.b, args.ldb, args.m, args.n, args.K, args.out,
                args.ldo, float(alpha), splits, stages)


def colsum_group_args(xs, outs) -> GroupArgs:
    return GroupArgs(x=("ptr", [x.data_ptr() for x in xs]), ld=("ll", [x.stride(0) for x in xs]),
                     m=("int", [x.shape[0] for x in xs]), n=("int", [x.shape[1] for x in xs]),
                     out=("ptr", [o.data_ptr() for o in outs]))


def colsum_grouped(args: GroupArgs) -> None:
    """out[p] += column sums of xs[p] for up to 4 matrices in one launch (the bias gradients of one layer)."""
    native.call("dtc_colsum_grouped", args.count, args.x, args.ld, args.m, args.n, args.out)


def cast_bf16(src_f32, dst_bf16):
    native.call("dtc_cast_f32_to_bf16", src_f32, dst_bf16, src_f32.numel())
    return dst_bf16


def sumsq_partials(g, partials):
    native.call("dtc_sumsq_partials", g, g.numel(), partials)
    return partials


def clip_finalize(partials, n_partials: int, grad_scale: float, max_norm: float, scalars, step):
    native.call("dtc_clip_finalize", partials, n_partials, float(grad_scale), float(max_norm), scalars, step)


def adamw(p, m, v, g, p_bf16, scalars, step_ptr, lr, b1, b2, eps, wd, grad_scale: float = 1.0, zero_grad: bool = True,
          n: int | None = None):
    native.call("dtc_adamw", p, m, v, g, p_bf16, p.numel() if n is None else n, scalars, step_ptr, float(lr), float(b1),
                float(b2), float(eps), float(wd), float(grad_scale), int(zero_grad))
