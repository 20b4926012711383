"""GPT model top (component C4; reference: model/GPTModel.py:21-82).

Same four callable pieces as the reference — ``embed_forward(x) -> (h, mask)``, ``stage_forward(h, mask,
n_layers) -> h``, ``head_forward(h) -> logits`` and ``__call__(x) -> logits`` — so a pipeline stage can run any
contiguous slice of the network.  Parameters are NOT owned by the module (as in Flax): every method takes the
dict of physical parameter views of a :class:`~.param_tree.ParamStore`.

  * torch path (fp32, autograd): wte[x] + wpe[arange(T)] → dropout(p, ALWAYS on — quirk Q13) → additive causal mask
    0/−1e9 → L × TransformerBlock (a Python loop over the stacked (L, ...) weights replaces ``nn.scan``, K15) →
    LayerNorm → untied lm_head with bias.
  * native path (bf16 compute, sm_100a kernels): fused embedding+position+Philox-dropout kernel, no mask tensor
    (causal predicate inside flash attention), vocab-padded lm_head GEMM + in-place softmax-CE kernel that
    turns the bf16 logits into dlogits (no fp32 logits tensor), see ``native_*`` methods.
"""
from __future__ import annotations

import os

import torch

from ..ops import functional as Fn
from ..parallel import tp_ops
from ..parallel.mesh import DeviceMesh, single_device_mesh
from .param_tree import LocalShape
from .TransformerBlock import TransformerBlock, layer_norm


class GPTModel:
    def __init__(self, config, local: LocalShape | None = None, mesh: DeviceMesh | None = None):
        self.config = config
        self.local = local or LocalShape(config)
        self.mesh = mesh or single_device_mesh("cpu")
        self.block = TransformerBlock(config, self.local, self.mesh)

    # ================================================================== torch path
    def embed_forward(self, p: dict, x: torch.Tensor, dropout_generator: torch.Generator | None = None):
        """x: (batch, T) integer ids → (h (batch,T,D), additive causal mask (1,1,T,T))."""
        _, T = x.shape
        h = p["wte"][x.long()] + p["wpe"][torch.arange(T, device=x.device)][None]
        pd = self.config.dropout
        if pd > 0:  # deterministic=False in the reference: dropout is always active
            keep = torch.rand(h.shape, device=h.device, generator=dropout_generator) >= pd
            h = torch.where(keep, h / (1.0 - pd), torch.zeros_like(h))
        h = tp_ops.slice_rows(h, self.mesh.tp)   # sequence-parallel TP: this rank's sequences only (identity otherwise)
        mask = torch.where(torch.tril(torch.ones(T, T, device=x.device)) == 1, 0.0, -1e9)[None, None]
        return h, mask

    def stage_forward(self, p: dict, h: torch.Tensor, mask: torch.Tensor | None, n_layers: int | None = None):
        n = self.local.L if n_layers is None else n_layers
        for l in range(n):
            h = self.block(p, l, h, mask)
        return h

    def head_forward(self, p: dict, h: torch.Tensor) -> torch.Tensor:
        """Returns the logits of this rank's (unpadded) vocabulary shard: (batch, T, vocab_valid)."""
        h = layer_norm(h, p["lnf_g"], p["lnf_b"], self.config.layernorm_eps)
        h = tp_ops.copy_to_tp(h, self.mesh.tp)
        nv = self.local.vocab_valid
        return h @ p["wlm"][:, :nv] + p["blm"][:nv]

    def __call__(self, p: dict, x: torch.Tensor, dropout_generator=None) -> torch.Tensor:
        h, mask = self.embed_forward(p, x, dropout_generator)
        h = self.stage_forward(p, h, mask)
        return self.head_forward(p, h)

    def loss_rows(self, p: dict, logits: torch.Tensor, y: torch.Tensor) -> torch.Tensor:
        """Per-token softmax cross-entropy with integer labels (reference: train/create_train_step.py:32-34)."""
        flat = logits.reshape(-1, logits.shape[-1])
        return tp_ops.vocab_parallel_cross_entropy(flat, y.reshape(-1), self.local.vocab_start, self.mesh.tp)

    # ================================================================== native path
    def native_embed_forward(self, eng, ws) -> None:
        c = eng.c
        Fn.embed_fwd(ws.ids, c["wte"], c["wpe"], ws.T, self.config.dropout, eng.seed, eng.store.step,
                     stream_id=ws.stream_id, row_offset=ws.row_offset, out=ws.h[0])

    def native_embed_backward(self, eng, ws, dh0: torch.Tensor) -> None:
        g = eng.g
        Fn.embed_bwd(ws.ids, dh0, g["wte"], g["wpe"], ws.T, self.config.dropout, eng.seed, eng.store.step,
                     stream_id=ws.stream_id, row_offset=ws.row_offset)

    def native_stage_forward(self, eng, ws) -> None:
        for l in range(self.local.L):
            self.block.native_forward(eng, ws, l)

    def native_stage_backward(self, eng, ws, dh_top: torch.Tensor) -> torch.Tensor:
        """Backward through the local layers; returns the buffer holding dL/dh[0]."""
        bufs = [ws.dh_a, ws.dh_b]
        cur = dh_top
        for l in reversed(range(self.local.L)):
            nxt = bufs[0] if cur is not bufs[0] else bufs[1]
            # the side stream (parameter gradients of layer l+1) still reads the scratch buffers this layer rewrites
            eng.side_join()
            self.block.native_backward(eng, ws, l, cur, ws.dh_mid, nxt)
            if eng.has_grad_hooks:
                eng.side_join()
                eng.on_layer_grads_ready(l)
            cur = nxt
        return cur

    def native_head_forward_backward(self, eng, ws, gscale: float) -> torch.Tensor:
        """LayerNorm → lm_head → softmax-CE (+ its gradient) → lm_head / LayerNorm backward.
        Writes the per-row losses to ws.loss_rows and returns the buffer holding dL/dh[L]."""
        c, g, lo, tp = eng.c, eng.g, self.local, self.mesh.tp
        L, eps = lo.L, self.config.layernorm_eps
        nv = lo.vocab_valid
        Fn.layernorm_fwd(ws.h[L], c["lnf_g"], c["lnf_b"], eps, out=ws.lnf, mean=ws.meanf, rstd=ws.rstdf)
        # Opt-in (DTC_CE_FUSED_STATS=1): the lm_head GEMM leaves per-part softmax statistics behind and the loss kernel makes
        # ONE pass over the logits.  Measured on B200: the extra epilogue work (rounding, masking, exp2 on 206 M elements in
        # the 8 epilogue warps) costs more than the saved 412 MB read: 4.47 vs 4.41 ms per step — so it is not the default.
        fused_stats = tp.size == 1 and ws.M >= 256 and os.environ.get("DTC_CE_FUSED_STATS", "0") == "1"
        if fused_stats:
            Fn.gemm(ws.lnf, c["wlm"], b_mn=True, bias=c["blm"], out=ws.logits, epi="store_stats", block_n=256,
                    stats=ws.ce_parts, n_valid=nv)
            Fn.ce_bwd_from_parts(ws.logits, ws.labels, ws.ce_parts, ws.ce_n_parts, nv, gscale, loss_rows=ws.loss_rows)
        elif tp.size == 1:
            Fn.gemm(ws.lnf, c["wlm"], b_mn=True, bias=c["blm"], out=ws.logits)
            Fn.ce_fwd_bwd(ws.logits, ws.labels, nv, gscale, loss_rows=ws.loss_rows)
        else:
            Fn.gemm(ws.lnf, c["wlm"], b_mn=True, bias=c["blm"], out=ws.logits)
            Fn.ce_stats(ws.logits, ws.labels, nv, lo.vocab_start, ws.ce_stats_local)
            tp.all_gather_into(ws.ce_stats_all, ws.ce_stats_local)
            Fn.ce_combine(ws.ce_stats_all, tp.size, ws.M, ws.lse, ws.loss_rows)
            Fn.ce_grad(ws.logits, ws.labels, ws.lse, nv, lo.vocab_start, gscale)
        # logits now hold dlogits (pad columns are zero)
        def head_param_grads():
            Fn.colsum(ws.logits, g["blm"])
            Fn.gemm(ws.lnf, ws.logits, a_mn=True, b_mn=True, epi="atomic_f32", out=g["wlm"], splits=1)

        eng.side_work(head_param_grads)
        # dL/d(lnf out) = dlogits · W_lmᵀ: [M, D] output (M/256 x D/256 pair tiles — 32 for the reference shapes, 4 for one
        # sequence per rank) over a reduction of V = 50304: split K so that every CTA pair has a slice, accumulate in fp32
        splits = max(1, 74 // max(1, -(-ws.M // 256) * -(-self.config.d_model // 256)))
        if tp.size == 1 and splits > 1:
            ws.dln_f32.zero_()
            Fn.gemm(ws.logits, c["wlm"], epi="atomic_f32", out=ws.dln_f32, splits=min(splits, 32))
            Fn.cast_bf16(ws.dln_f32, ws.dln)
            ws.dln_cur = ws.dln
        else:
            ws.dln_cur = eng.tp_reduced_gemm(ws.logits, c["wlm"], out=ws.dln)
        Fn.layernorm_bwd(ws.dln_cur, ws.h[L], ws.meanf, ws.rstdf, c["lnf_g"], g["lnf_g"], g["lnf_b"], dres=None, out=ws.dh_a)
        if eng.has_grad_hooks:
            eng.side_join()
            eng.on_head_grads_ready()
        return ws.dh_a
