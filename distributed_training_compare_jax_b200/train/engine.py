"""Stage engines: run forward / backward of the locally owned slice of the model for one microbatch.

Two interchangeable implementations behind the same interface (``set_batch`` / ``forward`` / ``backward``):

* :class:`NativeStageEngine` — the product path on B200: explicit (autograd-free) forward and backward over
  preallocated workspaces, every op a hand-written sm_100a kernel from ``libdtc_b200.so``; nothing is
  allocated after construction, so a whole training step can be captured into one CUDA graph.
* :class:`TorchStageEngine` — fp32 PyTorch autograd over the same parameter store (oracle T0, CPU / gloo
  plumbing path, and the "library kernels only" stand-in baseline when run on CUDA).

The strategies compose them: DP = one stage + gradient reduction, TP = collectives inside the stage,
PP = one stage per rank driven by a schedule (``parallel/pp.py``).  Replaces the reference's jitted
``value_and_grad(loss_fn)`` (train/create_train_step.py:28-50) and the per-tick body of the pmapped pipeline
step (train/create_train_step.py:90-168).
"""
from __future__ import annotations

import numpy as np
import os

import torch

from ..config.schema import ModelConfig
from ..models.GPTModel import GPTModel
from ..models.param_tree import LocalShape, ParamStore
from ..ops import functional as Fn
from ..parallel.mesh import DeviceMesh

BF16, F32, I32 = torch.bfloat16, torch.float32, torch.int32


class Workspace:
    """Activations saved for the backward pass + scratch, for one microbatch of ``B x T`` tokens."""

    def __init__(self, cfg: ModelConfig, local: LocalShape, B: int, T: int, device, tp_size: int, stream_id: int = 0):
        self.B, self.T, self.M = B, T, B * T
        self.stream_id, self.row_offset = stream_id, 0
        M, D, L = self.M, cfg.d_model, local.L
        Dl, Fl, Hl, Vl = local.d_local, local.ff_local, local.heads_local, local.vocab_local

        def e(*shape, dtype=BF16):
            return torch.empty(shape, device=device, dtype=dtype)

        self.ids = torch.zeros(M, device=device, dtype=I32)
        self.labels = torch.zeros(M, device=device, dtype=I32)
        self.h = [e(M, D) for _ in range(L + 1)]
        self.ln1, self.ln2 = e(L, M, D), e(L, M, D)
        self.mean1, self.rstd1 = e(L, M, dtype=F32), e(L, M, dtype=F32)
        self.mean2, self.rstd2 = e(L, M, dtype=F32), e(L, M, dtype=F32)
        self.qkv, self.attn = e(L, M, 3 * Dl), e(L, M, Dl)
        self.lse2 = e(L, B, Hl, T, dtype=F32)
        self.hmid = e(L, M, D)
        self.f1, self.g = e(L, M, Fl), e(L, M, Fl)
        # backward scratch (shared by all layers)
        self.dh_a, self.dh_b, self.dh_mid, self.dln = e(M, D), e(M, D), e(M, D), e(M, D)
        self.dqkv, self.dattn, self.df1 = e(M, 3 * Dl), e(M, Dl), e(M, Fl)
        # zero on entry of every attention backward and zero again when it returns (Fn.flash_attn_bwd(dq_clean=True))
        self.dq_accum, self.delta = torch.zeros(M, Dl, device=device, dtype=F32), e(B, Hl, T, dtype=F32)
        if local.has_head:
            self.lnf, self.meanf, self.rstdf = e(M, D), e(M, dtype=F32), e(M, dtype=F32)
            self.logits = e(M, Vl)
            self.dln_f32 = e(M, D, dtype=F32)   # split-K accumulator of the lm_head dgrad (few output tiles, K = vocab)
            self.ce_n_parts = 2 * (-(-Vl // 256))   # softmax statistics per 128-column part, from the lm_head GEMM epilogue
            self.ce_parts = e(M, self.ce_n_parts, 2, dtype=F32)
            self.loss_rows, self.lse = e(M, dtype=F32), e(M, dtype=F32)
            if tp_size > 1:
                self.ce_stats_local, self.ce_stats_all = e(M, 3, dtype=F32), e(tp_size, M, 3, dtype=F32)


class StageEngineBase:
    backend = "base"

    def __init__(self, cfg: ModelConfig, local: LocalShape, mesh: DeviceMesh, store: ParamStore, micro_batch: int,
                 seq_len: int, n_micro: int, global_tokens: int, seed: int):
        self.cfg, self.local, self.mesh, self.store = cfg, local, mesh, store
        self.micro_batch, self.seq_len, self.n_micro = micro_batch, seq_len, n_micro
        self.global_tokens = global_tokens  # tokens the mean loss is taken over (all DP ranks, all microbatches)
        self.seed = seed
        self.model = GPTModel(cfg, local, mesh)
        self.device = store.device
        self.loss_sum = torch.zeros(1, device=self.device, dtype=F32)  # sum over local rows / global_tokens
        self.has_grad_hooks = False  # set by the bucketed DP reducer

    # hooks used by the data-parallel engine to overlap gradient reduction with the backward pass
    def on_layer_grads_ready(self, l: int) -> None:
        pass

    def on_head_grads_ready(self) -> None:
        pass

    def begin_step(self) -> None:
        self.loss_sum.zero_()


class NativeStageEngine(StageEngineBase):
    backend = "native"

    def __init__(self, *a, wgrad_splits: int = 0, overlap_wgrad: bool = True, wgrad_grouped: bool | None = None, **kw):
        super().__init__(*a, **kw)
        # Weight-gradient work (wgrad GEMMs + bias column sums) does not feed the backward chain: it runs on a side
        # stream concurrently with the dgrad / attention / LayerNorm chain.  Both streams use the light (two CTAs per SM)
        # configuration of the persistent GEMM, so a CTA of each stream fits on one SM.
        self.overlap_wgrad = overlap_wgrad and __import__('os').environ.get('DTC_WGRAD_OVERLAP', '1') != '0'
        self.bwd_gemm_impl = int(__import__('os').environ.get('DTC_BWD_GEMM_IMPL', '0'))
        # One grouped wgrad launch + one grouped bias-grad launch per layer (default; measured on B200: 5.43 -> 5.21 ms per
        # step on one GPU, profiles/r2/).  DTC_WGRAD_GROUPED=0 restores the four split-K launches + four column sums.
        # The grouped kernel needs every dY width to be a multiple of 64 columns.
        if wgrad_grouped is None:
            wgrad_grouped = __import__('os').environ.get('DTC_WGRAD_GROUPED', '1') != '0'
        lo = self.local
        self.wgrad_grouped = bool(wgrad_grouped and self.cfg.d_model % 64 == 0 and lo.ff_local % 64 == 0
                                  and (3 * lo.d_local) % 64 == 0 and lo.d_local % 8 == 0)
        self.wgrad_pair = __import__('os').environ.get('DTC_WGRAD_PAIR', '1') != '0'   # CTA-pair grouped wgrad kernel
        self.side = None  # side stream of the workspace currently being processed (set in forward / backward)
        self._forked: set[int] = set()  # side streams that carry work not yet joined into the main stream
        if self.device.type != "cuda":
            raise RuntimeError("the native engine needs a CUDA device (sm_100a)")
        if self.seq_len % 128:
            raise ValueError("native flash attention needs seq_len % 128 == 0")
        self.c, self.g = self.store.c, self.store.g
        M = self.micro_batch * self.seq_len
        # split-K factor for the wgrad GEMMs (reduction over tokens): aim for >= 4 k-blocks of 64 per split
        self.wgrad_splits = wgrad_splits or max(1, min(8, M // 512))
        dp = self.mesh.dp
        self.ws = []
        for i in range(self.n_micro):
            w = Workspace(self.cfg, self.local, self.micro_batch, self.seq_len, self.device, self.mesh.tp.size,
                          stream_id=0)
            # global row index of local row 0: distinct dropout masks per DP rank / microbatch, identical across TP
            w.row_offset = (dp.rank * self.n_micro + i) * M
            w.side = torch.cuda.Stream(device=self.device) if self.overlap_wgrad else None
            if self.wgrad_grouped:
                # layer l's parameter gradients run while layer l-1's chain is already rewriting the scratch buffers:
                # two scratch sets (by layer parity) and a third dL/dh buffer keep their operands alive
                D, Dl, Fl = self.cfg.d_model, self.local.d_local, self.local.ff_local
                w.dh_c = torch.empty(M, D, device=self.device, dtype=BF16)
                w.scratch = [(w.df1, w.dqkv, w.dh_mid),
                             (torch.empty(M, Fl, device=self.device, dtype=BF16),
                              torch.empty(M, 3 * Dl, device=self.device, dtype=BF16),
                              torch.empty(M, D, device=self.device, dtype=BF16))]
                w.group_args = {}
            self.ws.append(w)

    # ---- building block shared by attention-out and fc2 (row-parallel Linear + bias + residual) ----
    def row_parallel_out(self, x, w, bias, residual, out) -> None:
        tp = self.mesh.tp
        if tp.size == 1 or tp.rank == 0:
            Fn.gemm(x, w, b_mn=True, bias=bias, epi="bias_residual", aux=residual, out=out)
        else:
            Fn.gemm(x, w, b_mn=True, out=out)
        tp.all_reduce(out)

    def side_work(self, fn, layer: int | None = None) -> None:
        """Run ``fn`` (kernel launches that only produce parameter gradients) on the side stream, after everything
        issued so far on the main stream.  ``layer``: set for the per-layer weight / bias gradients, which the grouped
        mode replaces by one launch per layer (``_stage_backward_grouped``)."""
        if layer is not None and self.wgrad_grouped:
            return
        if self.side is None:
            fn()
            return
        self.side.wait_stream(torch.cuda.current_stream())
        self._forked.add(id(self.side))
        with torch.cuda.stream(self.side):
            fn()

    def side_join(self) -> None:
        # Join only a side stream that was forked since its last join: waiting on a stream that carries no work of the
        # current capture is an error under stream capture ("dependency created on uncaptured work", the per-action
        # graphs of the pipeline executor) and a no-op otherwise.
        if self.side is not None and id(self.side) in self._forked:
            self._forked.discard(id(self.side))
            torch.cuda.current_stream().wait_stream(self.side)

    def backward_gemm_impl(self):
        # default policy: the persistent kernel in its light (<= 113 KB, 256 TMEM columns) configuration, two of which
        # share an SM — one from the backward chain, one from the parameter-gradient stream
        return Fn.gemm_impl(self.bwd_gemm_impl)

    def tp_reduced_gemm(self, a, w, out) -> torch.Tensor:
        """sum over TP ranks of a · wᵀ (dX of a column-parallel Linear, collectives X5 / X8); returns ``out``."""
        Fn.gemm(a, w, out=out)
        self.mesh.tp.all_reduce(out)
        return out

    def set_batch(self, mb: int, x: torch.Tensor | None, y: torch.Tensor | None) -> None:
        """x, y: int32 [micro_batch, T] (device or pinned host); copied into the static workspace buffers."""
        ws = self.ws[mb]
        if x is not None and self.local.has_embed:
            ws.ids.copy_(x.reshape(-1), non_blocking=True)
        if y is not None and self.local.has_head:
            ws.labels.copy_(y.reshape(-1), non_blocking=True)

    def input_buffer(self, mb: int) -> torch.Tensor:
        return self.ws[mb].h[0]

    def forward(self, mb: int) -> torch.Tensor | None:
        """Runs embed (first stage) + local layers (+ head, loss and head backward on the last stage).
        Returns the activation buffer to send downstream, or None on the last stage."""
        ws = self.ws[mb]
        self.side = ws.side
        if self.local.has_embed:
            self.model.native_embed_forward(self, ws)
        self.model.native_stage_forward(self, ws)
        if not self.local.has_head:
            return ws.h[self.local.L]
        ws.dh_top = self.model.native_head_forward_backward(self, ws, 1.0 / self.global_tokens)
        Fn.reduce_sum(ws.loss_rows, 1.0 / self.global_tokens, self.loss_sum, accumulate=True)
        self.side_join()
        return None

    def grad_input_buffer(self, mb: int) -> torch.Tensor:
        """Where an upstream gradient (dL/dh[L]) must be placed on a non-last stage."""
        return self.ws[mb].dh_b

    def backward(self, mb: int) -> torch.Tensor | None:
        ws = self.ws[mb]
        self.side = ws.side
        top = ws.dh_top if self.local.has_head else ws.dh_b
        with self.backward_gemm_impl():
            if self.wgrad_grouped:
                dh0 = self._stage_backward_grouped(ws, top)
            else:
                dh0 = self.model.native_stage_backward(self, ws, top)
        self.side_join()
        if self.local.has_embed:
            self.model.native_embed_backward(self, ws, dh0)
            return None
        return dh0


def _grouped_backward_impl(self, ws, top: torch.Tensor) -> torch.Tensor:
    """Backward through the local layers with ONE grouped weight-gradient launch and ONE grouped bias-gradient launch per
    layer on the side stream (instead of 4 split-K GEMMs + 4 column sums issued along the chain).  The launch of layer l
    runs concurrently with the chain of layer l-1; its operands stay alive because the scratch buffers alternate by
    layer parity and dL/dh rotates through three buffers; the chain of layer l-2 waits for it."""
    L, g = self.local.L, self.g
    main = torch.cuda.current_stream()
    ring = [ws.dh_a, ws.dh_b, ws.dh_c]
    i = next((k for k, b in enumerate(ring) if b.data_ptr() == top.data_ptr()), None)
    if i is None:            # gradient arrived in a foreign buffer: start the rotation next to it
        ring, i = [top, ws.dh_a, ws.dh_b], 0
    cur = top
    events: dict[int, torch.cuda.Event] = {}
    tiles = 0

    def release(layer: int) -> None:   # layer's parameter gradients are complete: buffers reusable, DP bucket reducible
        main.wait_event(events[layer])
        if self.has_grad_hooks:
            self.on_layer_grads_ready(layer)

    for l in reversed(range(L)):
        if l + 2 < L:
            release(l + 2)
        ws.df1, ws.dqkv, ws.dh_mid = ws.scratch[l & 1]
        i = (i + 1) % 3
        nxt = ring[i]
        self.model.block.native_backward(self, ws, l, cur, ws.dh_mid, nxt)
        key = (l, cur.data_ptr())
        if key not in ws.group_args:
            xs = [ws.g[l], ws.ln2[l], ws.attn[l], ws.ln1[l]]
            dys = [cur, ws.df1, ws.dh_mid, ws.dqkv]
            ws.group_args[key] = (Fn.colsum_group_args(dys, [g["b2"][l], g["b1"][l], g["bo"][l], g["bqkv"][l]]),
                                  Fn.wgrad_group_args(xs, dys, [g["w2"][l], g["w1"][l], g["wo"][l], g["wqkv"][l]]))
            tiles = sum(-(-x.shape[1] // 128) * -(-d.shape[1] // 128) for x, d in zip(xs, dys))
            ws.group_tiles256 = sum(-(-x.shape[1] // 256) * -(-d.shape[1] // 256) for x, d in zip(xs, dys))
        cs_args, wg_args = ws.group_args[key]
        splits = getattr(ws, "group_splits", None)
        if splits is None:
            if self.wgrad_pair:
                # 256 x 256 tiles on the 74 CTA pairs: slice the token reduction only while ALL slices still fit one wave.
                # (Alone, 3 slices of the small model's 48 tiles are faster — 27.4 vs 29.1 us, 1.95 waves — but in the step
                # this launch shares the GPU with the main stream: 3.79 ms with 3 slices, 3.74 ms with 1.)
                kb = max(1, self.micro_batch * self.seq_len // 64)
                splits = ws.group_splits = max(1, min(8, kb // 8, 74 // max(ws.group_tiles256, 1)))
                if os.environ.get("DTC_WGRAD_GROUP_SPLITS"):   # A/B override
                    splits = ws.group_splits = int(os.environ["DTC_WGRAD_GROUP_SPLITS"])
            else:                 # fill 2 CTAs per SM when the layer is small (tensor-parallel shards)
                splits = ws.group_splits = max(1, min(8, 296 // max(tiles, 1)))
        side = self.side if self.side is not None else main
        if side is not main:
            side.wait_stream(main)
        with torch.cuda.stream(side):
            Fn.colsum_grouped(cs_args)
            Fn.wgrad_grouped(wg_args, splits=splits, pair=self.wgrad_pair)
            ev = torch.cuda.Event()
            ev.record(side)
        events[l] = ev
        cur = nxt
    for l in (1, 0):
        if l < L:
            release(l)
    return cur


NativeStageEngine._stage_backward_grouped = _grouped_backward_impl


class TorchStageEngine(StageEngineBase):
    """fp32 autograd engine over the same ParamStore (CPU oracle / gloo path / cuBLAS stand-in baseline)."""
    backend = "torch"

    def __init__(self, *a, seq_parallel: bool = False, **kw):
        super().__init__(*a, **kw)
        # sequence-parallel TP (the fp32 counterpart of parallel/tp_sp.py): same conditions as the native engine
        tp = self.mesh.tp
        self.seq_parallel = bool(seq_parallel and tp.size > 1 and self.local.has_embed and self.local.has_head
                                 and self.n_micro == 1 and self.micro_batch % tp.size == 0)
        tp.seq_parallel = self.seq_parallel
        st = self.store
        self._partial_grads = st.flat_grad[st.offsets["ln1_g"][0]:st.rep_numel] if self.seq_parallel else None
        self.params = {}
        for name in self.store.names():
            p = self.store.p[name]
            p.requires_grad_(True)
            p.grad = self.store.g[name]
            self.params[name] = p
        self._x = [None] * self.n_micro
        self._y = [None] * self.n_micro
        self._in = [None] * self.n_micro
        self._out = [None] * self.n_micro
        self._loss = [None] * self.n_micro
        self._gin = [None] * self.n_micro
        self._step_count = 0

    def set_batch(self, mb, x, y) -> None:
        if x is not None:
            self._x[mb] = torch.as_tensor(np.asarray(x) if not torch.is_tensor(x) else x).to(self.device).long()
        if y is not None:
            self._y[mb] = torch.as_tensor(np.asarray(y) if not torch.is_tensor(y) else y).to(self.device).long()

    def begin_step(self) -> None:
        super().begin_step()
        self._step_count += 1

    def input_buffer(self, mb):
        if self._in[mb] is None:
            self._in[mb] = torch.zeros(self.micro_batch, self.seq_len, self.cfg.d_model, device=self.device)
        return self._in[mb]

    def forward(self, mb):
        m, p = self.model, self.params
        T = self.seq_len
        if self.local.has_embed:
            gen = torch.Generator(device=self.device)
            gen.manual_seed(self.seed * 7919 + (self._step_count * 64 + self.mesh.dp.rank) * 64 + mb)
            h, mask = m.embed_forward(p, self._x[mb], gen)
        else:
            h = self.input_buffer(mb).detach().requires_grad_(True)
            self._in[mb] = h
            mask = torch.where(torch.tril(torch.ones(T, T, device=self.device)) == 1, 0.0, -1e9)[None, None]
        h = m.stage_forward(p, h, mask)
        if not self.local.has_head:
            self._out[mb] = h
            return h.detach()
        logits = m.head_forward(p, h)
        rows = m.loss_rows(p, logits, self._y[mb])
        self._loss[mb] = rows.sum() / self.global_tokens
        self.loss_sum += self._loss[mb].detach()
        return None

    def grad_input_buffer(self, mb):
        if self._gin[mb] is None:
            self._gin[mb] = torch.zeros(self.micro_batch, self.seq_len, self.cfg.d_model, device=self.device)
        return self._gin[mb]

    def backward(self, mb):
        if self.local.has_head:
            self._loss[mb].backward()
        else:
            self._out[mb].backward(self.grad_input_buffer(mb))
        self._loss[mb], self._out[mb] = None, None
        if self.seq_parallel:   # LayerNorm / row-bias gradients come from the owned rows only: sum over the TP group
            self.mesh.tp.all_reduce(self._partial_grads)
        if self.local.has_embed:
            return None
        g = self._in[mb].grad
        return g


def make_engine(backend: str, *args, **kw) -> StageEngineBase:
    if backend == "native":
        kw.pop("seq_parallel", None)   # the native sequence-parallel engine is parallel/tp_sp.py (chosen in build_state)
        return NativeStageEngine(*args, **kw)
    if backend == "torch":
        kw.pop("wgrad_splits", None)
        return TorchStageEngine(*args, **kw)
    raise ValueError(f"unknown backend {backend}")
