"""Sequence-parallel tensor parallelism with fused GEMM→reduce-scatter and all-gather→GEMM (components R3 / R4 of
SURVEY.md §8; collectives X3 / X4 / X5 / X8 of §2.5 — the all-reduces GSPMD derives from the reference's
``with_sharding_constraint(..., replicated)`` after the row-parallel out_proj / fc2 matmuls,
model/CausalSelfAttention.py:49-50 and model/MLP.py:23-24, and their transposes in the backward pass).

The plain TP path (``train/engine.py``) keeps the residual stream replicated and all-reduces every row-parallel output
with NCCL: 4 all-reduces of [B·T, D] per layer and step, each a separate pass over the data, plus LayerNorm / residual
work repeated on every rank.  Here every all-reduce is split into its two halves and each half is folded into the
kernel next to it (``tp_seq_parallel: true``):

* reduce-scatter = the EPILOGUE of the producing GEMM (``dtc_gemm_bf16_scatter``): each 128-row output tile is stored
  over NVLink into the owner rank's partial-sum slot while later tiles are still in the tensor cores;
* the owner (M/n rows of the residual stream per rank) sums the n slots in fp32 inside its LayerNorm kernel
  (``dtc_sp_ln_fwd`` / ``dtc_sp_ln_bwd``), which also adds bias + residual (forward) or the residual gradient
  (backward) — work that is now divided by n;
* all-gather = the STORES of that same LayerNorm kernel: its output rows go to every rank's operand buffer through the
  NVSwitch multicast address (``multimem.st``), so the following column-parallel GEMM (q/k/v, fc1, lm_head, and the
  dgrad / wgrad GEMMs of the backward pass) finds its full [B·T, D] operand in local memory.

Between a producer and its consumer on another rank there is one flag barrier (tiny kernel, no host sync), so the whole
step stays capturable in one CUDA graph.  Numerically the path differs from plain TP only by summing the partial
results in fp32 before the single rounding to bf16 (NCCL rounds after every hop).

Gradients of replicated leaves that are computed from owned rows only (LayerNorm scale / bias, out_proj / fc2 bias) are
partial sums: they live in ONE contiguous range of the flat gradient buffer (``ln1_g … lnf_b``) and are summed over the
TP group by a single all-reduce after the backward pass.
"""
from __future__ import annotations

import ctypes as C

import torch

from ..ops import functional as Fn
from ..train.engine import NativeStageEngine
from .symm import FlagBarrier, SymmBuffer

BF16, F32 = torch.bfloat16, torch.float32


class _LocalArena:
    """Stand-in for :class:`SymmBuffer` when the TP group has one rank (unit tests of the data flow on one GPU)."""

    def __init__(self, numel: int, device):
        self.tensor = torch.zeros(numel, device=device, dtype=BF16)
        self.ptrs = [self.tensor.data_ptr()]
        self.mc_ptr = 0
        self.rank, self.world = 0, 1


class SeqParallelState:
    """Buffers and cross-rank plumbing of the sequence-parallel path for one microbatch of M = B·T rows."""

    def __init__(self, engine, mesh, use_multicast: bool = True, fold_barriers: bool = False):
        tp = mesh.tp
        cfg, local = engine.cfg, engine.local
        self.n, self.rank = tp.size, tp.rank
        self.M = engine.micro_batch * engine.seq_len
        self.D, self.L = cfg.d_model, local.L
        n, M, D, L = self.n, self.M, self.D, self.L
        if M % (n * 128):
            raise ValueError(f"sequence-parallel TP needs B·T = {M} to be a multiple of 128·tp = {128 * n}")
        self.Ms = M // n
        if self.Ms % engine.seq_len:
            raise ValueError(f"sequence-parallel TP needs whole sequences per rank: batch {engine.micro_batch} % tp {n} != 0")
        self.row0 = self.rank * self.Ms
        dev = engine.device
        # ---- symmetric arena: gathered operands (saved for wgrad), gathered gradients, partial-sum slots
        n_full = 2 * L + 1 + 3 + 1
        self.arena = SymmBuffer(n_full * M * D, BF16, dev, tp.group) if n > 1 else _LocalArena(n_full * M * D, dev)
        self.use_mc = bool(use_multicast and self.arena.mc_ptr)
        full = self.arena.tensor.view(n_full, M, D)
        self.xg1, self.xg2 = full[:L], full[L:2 * L]          # LayerNorm outputs of every layer, all rows
        self.xgf = full[2 * L]
        self.dA, self.dB, self.dMID = full[2 * L + 1], full[2 * L + 2], full[2 * L + 3]
        self.part = full[2 * L + 4]                            # viewed by the kernels as [n, Ms, D]
        self.barrier_flags = FlagBarrier(dev, tp.group) if n > 1 else None
        if self.barrier_flags is not None:
            self.barrier_flags.step_counter = engine.store.step   # (jitter stress test: a device int that changes per step)
        self._slot = 0
        self._ptr_cache: dict[int, tuple] = {}
        # ---- local: owned rows of the residual stream and the LayerNorm statistics
        Ms = self.Ms

        def e(*shape, dtype=BF16):
            return torch.empty(shape, device=dev, dtype=dtype)

        self.hs = [e(Ms, D) for _ in range(L + 1)]
        self.hmids = [e(Ms, D) for _ in range(L)]
        self.mean1, self.rstd1 = e(L, Ms, dtype=F32), e(L, Ms, dtype=F32)
        self.mean2, self.rstd2 = e(L, Ms, dtype=F32), e(L, Ms, dtype=F32)
        self.meanf, self.rstdf = e(Ms, dtype=F32), e(Ms, dtype=F32)
        self.part_ptrs = self._peer_array(self.part)
        # small fp32 exchanges (cross-entropy statistics, partial gradients of replicated leaves): one slot per rank on
        # every rank, filled by dtc_symm_push_f32 — no NCCL call is left inside the step graph
        self.ce_n = M * 3
        if n > 1:
            st = engine.store
            self.pg_off, self.pg_n = st.offsets["ln1_g"][0], st.rep_numel - st.offsets["ln1_g"][0]
            self.small = SymmBuffer(n * (self.ce_n + self.pg_n), F32, dev, tp.group)
            self.ce_slots = self.small.tensor[:n * self.ce_n].view(n, M, 3)
            self.pg_slots = self.small.tensor[n * self.ce_n:]
        # Experimental (DTC_SP_FOLD=1, not yet run on hardware): no barrier kernels between producers and consumers — the
        # last CTA of a scatter GEMM / LayerNorm kernel publishes a sequence number to every rank, the LayerNorm kernels wait
        # in their prologue and a one-CTA wait kernel sits in front of the stock GEMMs that read a gathered operand.
        self.fold = bool(fold_barriers)
        if self.fold:
            if n > 1:
                self.sig = SymmBuffer(16, torch.int32, dev, tp.group)
                self.sig_ptrs = self.sig.ptr_array
            else:
                self.sig = torch.zeros(16, device=dev, dtype=torch.int32)
                self.sig_ptrs = (C.c_ulonglong * 1)(self.sig.data_ptr())
            self.counters = torch.zeros(4, device=dev, dtype=torch.int32)

    # ---- addressing ----
    def _offset_bytes(self, view: torch.Tensor) -> int:
        return view.data_ptr() - self.arena.tensor.data_ptr()

    def _peer_array(self, view: torch.Tensor):
        off = self._offset_bytes(view)
        return (C.c_ulonglong * self.n)(*[p + off for p in self.arena.ptrs])

    def dest(self, view: torch.Tensor) -> tuple:
        """(multicast address or 0, ctypes array of every rank's address of ``view``, n) — cached per view."""
        key = view.data_ptr()
        hit = self._ptr_cache.get(key)
        if hit is None:
            off = self._offset_bytes(view)
            hit = (self.arena.mc_ptr + off if self.use_mc else 0, self._peer_array(view), self.n)
            self._ptr_cache[key] = hit
        return hit

    # ---- primitives ----
    def barrier(self) -> None:
        if self.barrier_flags is not None:
            self.barrier_flags(self._slot)
            self._slot = (self._slot + 1) % FlagBarrier.N_SLOTS

    def scattered(self) -> None:
        """After a scatter GEMM, before the kernel that reduces the partial-sum slots."""
        if not self.fold:
            self.barrier()          # folded mode: the GEMM signals, the LayerNorm kernel waits in its prologue

    def pushed(self) -> None:
        """After a LayerNorm kernel pushed its rows to every rank, before the GEMMs that read the gathered buffer."""
        if self.fold:
            Fn.sp_wait(self.sig_ptrs, self.counters, self.rank, self.n)
        else:
            self.barrier()

    def gemm_scatter(self, a: torch.Tensor, w: torch.Tensor, b_mn: bool) -> None:
        if self.fold:
            Fn.gemm_scatter_signal(a, w, self.part_ptrs, self.n, self.rank, self.D, self.sig_ptrs, self.counters, b_mn=b_mn)
        else:
            Fn.gemm_scatter(a, w, self.part_ptrs, self.n, self.rank, self.D, b_mn=b_mn)

    def ln_fwd(self, res, bias, reduce: bool, gamma, beta, hsum, mean, rstd, out, eps: float) -> None:
        mc, peers, n = self.dest(out)
        part, n_part = (self.part, self.n) if reduce else (None, 0)
        if self.fold:
            Fn.sp_ln_fwd_sync(res, bias, part, n_part, gamma, beta, hsum, mean, rstd, mc, peers, n, self.row0, eps,
                              self.sig_ptrs, self.counters, self.rank, self.n, wait_in=reduce, signal_out=True)
        else:
            Fn.sp_ln_fwd(res, bias, part, n_part, gamma, beta, hsum, mean, rstd, mc, peers, n, self.row0, eps)

    def ln_bwd(self, x, mean, rstd, gamma, dres, dgamma, dbeta, out) -> None:
        mc, peers, n = self.dest(out)
        if self.fold:
            Fn.sp_ln_bwd_sync(self.part, self.n, x, mean, rstd, gamma, dres, dgamma, dbeta, mc, peers, n, self.row0,
                              self.sig_ptrs, self.counters, self.rank, self.n, wait_in=True, signal_out=True)
        else:
            Fn.sp_ln_bwd(self.part, self.n, x, mean, rstd, gamma, dres, dgamma, dbeta, mc, peers, n, self.row0)

    def all_gather_small(self, local: torch.Tensor) -> torch.Tensor:
        """[M, 3] fp32 statistics of every rank -> [n, M, 3] (push into slot `rank` everywhere + flag barrier)."""
        from ..ops import native
        native.call("dtc_symm_push_f32", local, self.small.mc_ptr if self.use_mc else 0, self.small.ptr_array, self.n,
                    self.rank * self.ce_n, self.ce_n)
        self.barrier()
        return self.ce_slots

    def all_reduce_partial_grads(self, flat_grad: torch.Tensor) -> None:
        """Sum over the TP group of the partial-gradient range (LayerNorm scale / bias, out_proj / fc2 bias), in place."""
        from ..ops import native
        mine = flat_grad[self.pg_off:self.pg_off + self.pg_n]
        native.call("dtc_symm_push_f32", mine, self.small.mc_ptr if self.use_mc else 0, self.small.ptr_array, self.n,
                    self.n * self.ce_n + self.rank * self.pg_n, self.pg_n)
        self.barrier()
        native.call("dtc_sum_slots_f32", self.pg_slots, self.pg_n, self.n, mine, self.pg_n, 0)

    def own(self, full: torch.Tensor) -> torch.Tensor:
        return full[self.row0:self.row0 + self.Ms]


class SeqParallelEngine(NativeStageEngine):
    """Native stage engine whose residual stream is sharded by rows over the TP group (see the module docstring).
    Same interface as :class:`NativeStageEngine`; needs the whole model on the stage (no pipeline split) and one
    microbatch."""

    defer_grad_reduction = True   # parallel/dp.py: all-reduce the DP gradient after backward, not bucket by bucket

    def __init__(self, *a, use_multicast: bool = True, fold_barriers: bool | None = None, **kw):
        super().__init__(*a, **kw)
        if fold_barriers is None:
            fold_barriers = __import__("os").environ.get("DTC_SP_FOLD") == "1"
        if not (self.local.has_embed and self.local.has_head) or self.n_micro != 1 or self.local.L < 1:
            raise ValueError("sequence-parallel TP runs the whole model on one stage with one microbatch")
        ws = self.ws[0]
        for name in ("h", "ln1", "ln2", "hmid", "mean1", "rstd1", "mean2", "rstd2", "dh_a", "dh_b", "dh_mid", "dln", "lnf",
                     "meanf", "rstdf"):
            if hasattr(ws, name):
                delattr(ws, name)   # replaced by the row-sharded / gathered buffers of SeqParallelState
        self.sp = SeqParallelState(self, self.mesh, use_multicast, fold_barriers)
        st = self.store
        self._partial_grads = st.flat_grad[st.offsets["ln1_g"][0]:st.rep_numel]   # ln1_g ln1_b bo ln2_g ln2_b b2 lnf_g lnf_b

    def input_buffer(self, mb: int):
        raise RuntimeError("sequence-parallel TP has no pipeline input")

    def forward(self, mb: int):
        ws, sp, c, g, lo = self.ws[mb], self.sp, self.c, self.g, self.local
        self.side = ws.side
        cfg, tp = self.cfg, self.mesh.tp
        L, eps, n = lo.L, cfg.layernorm_eps, sp.n
        Hl, hd = lo.heads_local, cfg.head_dim
        sp._slot = 0
        sp.barrier()   # every rank has left the previous step: gathered buffers and partial slots may be rewritten
        Fn.embed_fwd(ws.ids[sp.row0:sp.row0 + sp.Ms], c["wte"], c["wpe"], ws.T, cfg.dropout, self.seed, self.store.step,
                     stream_id=ws.stream_id, row_offset=ws.row_offset + sp.row0, out=sp.hs[0])
        for l in range(L):
            if l == 0:
                sp.ln_fwd(sp.hs[0], None, False, c["ln1_g"][0], c["ln1_b"][0], None, sp.mean1[0], sp.rstd1[0], sp.xg1[0], eps)
            else:   # h[l] = hmid[l-1] + b2 + Σ fc2 partials
                sp.ln_fwd(sp.hmids[l - 1], c["b2"][l - 1], True, c["ln1_g"][l], c["ln1_b"][l], sp.hs[l], sp.mean1[l],
                          sp.rstd1[l], sp.xg1[l], eps)
            sp.pushed()
            Fn.gemm(sp.xg1[l], c["wqkv"][l], b_mn=True, bias=c["bqkv"][l], out=ws.qkv[l])
            Fn.flash_attn_fwd(ws.qkv[l], ws.B, Hl, ws.T, hd, out=ws.attn[l], lse2=ws.lse2[l])
            sp.gemm_scatter(ws.attn[l], c["wo"][l], b_mn=True)
            sp.scattered()
            # hmid[l] = h[l] + bo + Σ out_proj partials
            sp.ln_fwd(sp.hs[l], c["bo"][l], True, c["ln2_g"][l], c["ln2_b"][l], sp.hmids[l], sp.mean2[l], sp.rstd2[l],
                      sp.xg2[l], eps)
            sp.pushed()
            Fn.gemm(sp.xg2[l], c["w1"][l], b_mn=True, bias=c["b1"][l], epi="bias_gelu", out=ws.f1[l], out2=ws.g[l])
            sp.gemm_scatter(ws.g[l], c["w2"][l], b_mn=True)
            sp.scattered()
        # ---- head: final LayerNorm (gathered) → vocab-parallel lm_head → softmax-CE and its gradient
        sp.ln_fwd(sp.hmids[L - 1], c["b2"][L - 1], True, c["lnf_g"], c["lnf_b"], sp.hs[L], sp.meanf, sp.rstdf, sp.xgf, eps)
        sp.pushed()
        nv, gscale = lo.vocab_valid, 1.0 / self.global_tokens
        Fn.gemm(sp.xgf, c["wlm"], b_mn=True, bias=c["blm"], out=ws.logits)
        if n == 1:
            Fn.ce_fwd_bwd(ws.logits, ws.labels, nv, gscale, loss_rows=ws.loss_rows)
        else:
            Fn.ce_stats(ws.logits, ws.labels, nv, lo.vocab_start, ws.ce_stats_local)
            Fn.ce_combine(sp.all_gather_small(ws.ce_stats_local), n, ws.M, ws.lse, ws.loss_rows)
            Fn.ce_grad(ws.logits, ws.labels, ws.lse, nv, lo.vocab_start, gscale)

        def head_param_grads():   # logits now hold dlogits (pad columns are zero)
            Fn.colsum(ws.logits, g["blm"])
            Fn.gemm(sp.xgf, ws.logits, a_mn=True, b_mn=True, epi="atomic_f32", out=g["wlm"], splits=1)

        self.side_work(head_param_grads)
        sp.gemm_scatter(ws.logits, c["wlm"], b_mn=False)
        sp.scattered()
        sp.ln_bwd(sp.hs[L], sp.meanf, sp.rstdf, c["lnf_g"], None, g["lnf_g"], g["lnf_b"], sp.dA)
        sp.pushed()
        Fn.reduce_sum(ws.loss_rows, gscale, self.loss_sum, accumulate=True)
        self.side_join()
        return None

    def grad_input_buffer(self, mb: int):
        raise RuntimeError("sequence-parallel TP has no pipeline input")

    def backward(self, mb: int):
        ws, sp, c, g, lo = self.ws[mb], self.sp, self.c, self.g, self.local
        self.side = ws.side
        cfg = self.cfg
        L, Hl, hd, ks = lo.L, lo.heads_local, cfg.head_dim, self.wgrad_splits
        cur, other = sp.dA, sp.dB      # cur holds dL/dh[l+1] for ALL rows on every rank
        with self.backward_gemm_impl():
            for l in reversed(range(L)):
                # the side stream (parameter gradients of layer l+1) still reads buffers this layer rewrites
                self.side_join()
                cur_own = sp.own(cur)

                # ---- MLP
                def fc2_param_grads(cur=cur, cur_own=cur_own, l=l):
                    Fn.colsum(cur_own, g["b2"][l])          # owned rows only: summed over the TP group after backward
                    Fn.gemm(ws.g[l], cur, a_mn=True, b_mn=True, epi="atomic_f32", out=g["w2"][l], splits=ks)

                def fc1_param_grads(l=l):
                    Fn.colsum(ws.df1, g["b1"][l])
                    Fn.gemm(sp.xg2[l], ws.df1, a_mn=True, b_mn=True, epi="atomic_f32", out=g["w1"][l], splits=ks)

                self.side_work(fc2_param_grads)
                Fn.gemm(cur, c["w2"][l], epi="dgelu", aux=ws.f1[l], out=ws.df1)
                self.side_work(fc1_param_grads)
                sp.gemm_scatter(ws.df1, c["w1"][l], b_mn=False)
                sp.scattered()
                sp.ln_bwd(sp.hmids[l], sp.mean2[l], sp.rstd2[l], c["ln2_g"][l], cur_own, g["ln2_g"][l], g["ln2_b"][l], sp.dMID)
                sp.pushed()

                # ---- attention
                mid_own = sp.own(sp.dMID)

                def out_param_grads(mid_own=mid_own, l=l):
                    Fn.colsum(mid_own, g["bo"][l])
                    Fn.gemm(ws.attn[l], sp.dMID, a_mn=True, b_mn=True, epi="atomic_f32", out=g["wo"][l], splits=ks)

                def qkv_param_grads(l=l):
                    Fn.colsum(ws.dqkv, g["bqkv"][l])
                    Fn.gemm(sp.xg1[l], ws.dqkv, a_mn=True, b_mn=True, epi="atomic_f32", out=g["wqkv"][l], splits=ks)

                self.side_work(out_param_grads)
                Fn.gemm(sp.dMID, c["wo"][l], out=ws.dattn)
                Fn.flash_attn_bwd(ws.qkv[l], ws.attn[l], ws.dattn, ws.lse2[l], ws.B, Hl, ws.T, hd, dqkv=ws.dqkv,
                                  dq_accum=ws.dq_accum, delta=ws.delta, dq_clean=True)
                self.side_work(qkv_param_grads)
                sp.gemm_scatter(ws.dqkv, c["wqkv"][l], b_mn=False)
                sp.scattered()
                sp.ln_bwd(sp.hs[l], sp.mean1[l], sp.rstd1[l], c["ln1_g"][l], mid_own, g["ln1_g"][l], g["ln1_b"][l], other)
                sp.pushed()
                cur, other = other, cur
        self.side_join()
        # dL/dh[0] is complete on every rank: the embedding tables are replicated, every rank scatters all rows
        Fn.embed_bwd(ws.ids, cur, g["wte"], g["wpe"], ws.T, cfg.dropout, self.seed, self.store.step,
                     stream_id=ws.stream_id, row_offset=ws.row_offset)
        if sp.n > 1:
            sp.all_reduce_partial_grads(self.store.flat_grad)
        return None
