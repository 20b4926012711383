"""Parameter tree: Flax-compatible naming on top of flat, kernel-friendly device buffers.

Logical view  — exactly the reference's Flax tree (model/GPTModel.py:30-72 via ``nn.scan`` → ``ScanScanBlock_0``):
    wte/embedding (V,D)  wpe/embedding (T,D)
    ScanScanBlock_0/TransformerBlock_0/{LayerNorm_0,LayerNorm_1}/{scale,bias} (L,D)
    .../CausalSelfAttention_0/{q_proj,k_proj,v_proj,out_proj}/{kernel (L,in,out), bias (L,out)}
    .../MLP_0/{fc1,fc2}/{kernel,bias}      LayerNorm_0/{scale,bias} (D,)      lm_head/{kernel (D,V), bias (V,)}
  kernels are (in, out) like Flax (y = x·W + b), layers stacked on axis 0, vocab unpadded (50258).

Physical view — what one rank holds: a handful of stacked tensors carved out of ONE flat fp32 master buffer
(plus same-layout flat gradient / Adam m / Adam v buffers and a bf16 compute copy), so that the optimizer,
gradient norm and DP collectives are single launches over contiguous memory:
    replicated group : wte wpe ln1_g ln1_b bo ln2_g ln2_b b2 lnf_g lnf_b
    sharded group    : wqkv (L,D,3·D/tp) bqkv wo (L,D/tp,D) w1 (L,D,F/tp) b1 w2 (L,F/tp,D) wlm (D,Vl) blm (Vl)
  q/k/v projections are fused along the output axis ([q|k|v] of the local heads); the vocabulary is padded to
  a multiple of 128 per tensor-parallel rank (Vl) — pad columns are zero and masked in the loss (quirk Q1).
The tensor-parallel split follows the reference rule table parallel/sharding.py:29-60 (column-parallel q/k/v/fc1,
row-parallel out_proj/fc2, vocab-parallel lm_head); a pipeline stage holds a contiguous layer range.
"""
from __future__ import annotations

import math
import zlib
from dataclasses import dataclass
from typing import Iterable

import numpy as np
import torch

from ..config.schema import ModelConfig

BLOCK = "ScanScanBlock_0/TransformerBlock_0"
ATTN = f"{BLOCK}/CausalSelfAttention_0"
MLP_ = f"{BLOCK}/MLP_0"
ALIGN = 64  # elements; keeps every tensor 256-byte (fp32) / 128-byte (bf16) aligned inside the flat buffers


def flax_param_shapes(cfg: ModelConfig, n_layers: int | None = None) -> dict[str, tuple[int, ...]]:
    """Path → shape of the reference's DP/TP ``params`` tree."""
    L = cfg.n_layers if n_layers is None else n_layers
    D, F, V, T = cfg.d_model, cfg.d_ff, cfg.vocab_size, cfg.max_seq_len
    s: dict[str, tuple[int, ...]] = {"wte/embedding": (V, D), "wpe/embedding": (T, D)}
    for ln in ("LayerNorm_0", "LayerNorm_1"):
        s[f"{BLOCK}/{ln}/scale"] = (L, D)
        s[f"{BLOCK}/{ln}/bias"] = (L, D)
    for proj in ("q_proj", "k_proj", "v_proj", "out_proj"):
        s[f"{ATTN}/{proj}/kernel"] = (L, D, D)
        s[f"{ATTN}/{proj}/bias"] = (L, D)
    s[f"{MLP_}/fc1/kernel"] = (L, D, F)
    s[f"{MLP_}/fc1/bias"] = (L, F)
    s[f"{MLP_}/fc2/kernel"] = (L, F, D)
    s[f"{MLP_}/fc2/bias"] = (L, D)
    s["LayerNorm_0/scale"] = (D,)
    s["LayerNorm_0/bias"] = (D,)
    s["lm_head/kernel"] = (D, V)
    s["lm_head/bias"] = (V,)
    return s


def count_params(cfg: ModelConfig) -> int:
    return sum(int(np.prod(s)) for s in flax_param_shapes(cfg).values())


@dataclass(frozen=True)
class LocalShape:
    """Which slice of the model one rank owns."""
    cfg: ModelConfig
    layer_start: int = 0
    layer_end: int = -1      # exclusive; -1 = n_layers
    has_embed: bool = True
    has_head: bool = True
    tp_rank: int = 0
    tp_size: int = 1

    @property
    def layers(self) -> range:
        return range(self.layer_start, self.cfg.n_layers if self.layer_end < 0 else self.layer_end)

    @property
    def L(self) -> int:
        return len(self.layers)

    @property
    def d_local(self) -> int:
        return self.cfg.d_model // self.tp_size

    @property
    def heads_local(self) -> int:
        return self.cfg.n_heads // self.tp_size

    @property
    def ff_local(self) -> int:
        return self.cfg.d_ff // self.tp_size

    @property
    def vocab_local(self) -> int:
        """Padded per-rank vocabulary width Vl (multiple of 128)."""
        per = -(-self.cfg.vocab_size // self.tp_size)
        return -(-per // 128) * 128

    @property
    def vocab_start(self) -> int:
        return self.tp_rank * self.vocab_local

    @property
    def vocab_valid(self) -> int:
        """Number of real (unpadded) vocabulary columns held by this rank."""
        return max(0, min(self.cfg.vocab_size, self.vocab_start + self.vocab_local) - self.vocab_start)

    def physical_shapes(self) -> tuple[list[tuple[str, tuple[int, ...]]], list[tuple[str, tuple[int, ...]]]]:
        c, L = self.cfg, self.L
        D, Dl, Fl, Vl = c.d_model, self.d_local, self.ff_local, self.vocab_local
        rep: list[tuple[str, tuple[int, ...]]] = []
        shard: list[tuple[str, tuple[int, ...]]] = []
        if self.has_embed:
            rep += [("wte", (c.vocab_size, D)), ("wpe", (c.max_seq_len, D))]
        if L > 0:
            rep += [("ln1_g", (L, D)), ("ln1_b", (L, D)), ("bo", (L, D)), ("ln2_g", (L, D)), ("ln2_b", (L, D)),
                    ("b2", (L, D))]
            shard += [("wqkv", (L, D, 3 * Dl)), ("bqkv", (L, 3 * Dl)), ("wo", (L, Dl, D)), ("w1", (L, D, Fl)),
                      ("b1", (L, Fl)), ("w2", (L, Fl, D))]
        if self.has_head:
            rep += [("lnf_g", (D,)), ("lnf_b", (D,))]
            shard += [("wlm", (D, Vl)), ("blm", (Vl,))]
        return rep, shard


class ParamStore:
    """Flat fp32 master / grad / Adam-m / Adam-v buffers (+ bf16 compute copy) with named views.

    Layout: [ replicated group | sharded group ], every tensor ALIGN-element aligned.  ``self.p[name]``,
    ``self.g[name]``, ``self.c[name]`` (compute copy) are views; ``flat_*`` are the whole buffers."""

    def __init__(self, local: LocalShape, device: torch.device | str = "cpu", with_compute_copy: bool = True,
                 with_optimizer_state: bool = True, flat_alloc=None, symm_group=None):
        self.local = local
        self.device = torch.device(device)
        rep, shard = local.physical_shapes()
        self.offsets: dict[str, tuple[int, tuple[int, ...]]] = {}
        off = 0
        for name, shape in rep:
            self.offsets[name] = (off, shape)
            off += -(-int(np.prod(shape)) // ALIGN) * ALIGN
        self.rep_numel = off
        for name, shape in shard:
            self.offsets[name] = (off, shape)
            off += -(-int(np.prod(shape)) // ALIGN) * ALIGN
        self.numel = off
        self.shard_numel = off - self.rep_numel
        alloc = flat_alloc or (lambda n, dtype: torch.zeros(n, device=self.device, dtype=dtype))
        self.flat_master = alloc(self.numel, torch.float32)
        self.flat_m = alloc(self.numel, torch.float32) if with_optimizer_state else None
        self.flat_v = alloc(self.numel, torch.float32) if with_optimizer_state else None
        self.symm = None
        if symm_group is not None:
            # gradient + bf16 compute copy in symmetric memory: peers (and the NVSwitch multicast address) can reach
            # them, which is what the fused reduce-scatter / sharded AdamW / all-gather kernels operate on
            from ..parallel.symm import SymmBuffer
            self.symm = {"grad": SymmBuffer(self.numel, torch.float32, self.device, symm_group),
                         "compute": SymmBuffer(self.numel, torch.bfloat16, self.device, symm_group)}
            self.flat_grad = self.symm["grad"].tensor
            self.flat_compute = self.symm["compute"].tensor
        else:
            self.flat_grad = alloc(self.numel, torch.float32)
            self.flat_compute = alloc(self.numel, torch.bfloat16) if with_compute_copy else None
        self.p = self._views(self.flat_master)
        self.g = self._views(self.flat_grad)
        self.c = self._views(self.flat_compute) if with_compute_copy else None
        self.step = torch.zeros(1, dtype=torch.int32, device=self.device)  # optimizer step count (device side)

    def _views(self, flat: torch.Tensor) -> dict[str, torch.Tensor]:
        return {n: flat[o:o + int(np.prod(s))].view(s) for n, (o, s) in self.offsets.items()}

    def views_of(self, flat: torch.Tensor) -> dict[str, torch.Tensor]:
        return self._views(flat)

    def names(self) -> Iterable[str]:
        return self.offsets.keys()

    def sync_compute_copy(self) -> None:
        if self.flat_compute is not None:
            self.flat_compute.copy_(self.flat_master)

    def zero_grad(self) -> None:
        self.flat_grad.zero_()


# ----------------------------------------------------------------------------------------------------------
# Initialisation (Flax defaults, SURVEY.md §2.6 "Init" column)
# ----------------------------------------------------------------------------------------------------------
def _gen(seed: int, device, *key) -> torch.Generator:
    h = zlib.crc32(repr(key).encode())
    g = torch.Generator(device=device)
    g.manual_seed((seed * 1000003 + h) % (2 ** 63 - 1))
    return g


def _lecun_normal(shape, fan_in: int, g: torch.Generator) -> torch.Tensor:
    std = math.sqrt(1.0 / fan_in) / 0.87962566103423978  # truncated-normal variance correction (jax)
    t = torch.empty(shape, dtype=torch.float32, device=g.device)
    torch.nn.init.trunc_normal_(t, mean=0.0, std=std, a=-2 * std, b=2 * std, generator=g)
    return t


def init_logical_leaf(cfg: ModelConfig, path: str, layer: int | None, seed: int, key_prefix=(),
                      device="cpu") -> torch.Tensor:
    """One logical (unsharded, unpadded) leaf — for stacked leaves, the slice of layer ``layer``.  The random
    stream depends on (seed, path, layer) and on the DEVICE TYPE the generator runs on: every rank of a job
    draws identical values, CPU and CUDA jobs draw different (equally distributed) ones."""
    D, F, V, T = cfg.d_model, cfg.d_ff, cfg.vocab_size, cfg.max_seq_len
    g = _gen(seed, device, *key_prefix, path, layer)
    if path == "wte/embedding":
        return torch.randn((V, D), generator=g, device=g.device) / math.sqrt(D)
    if path == "wpe/embedding":
        return torch.randn((T, D), generator=g, device=g.device) / math.sqrt(D)
    if path.endswith("/scale"):
        return torch.ones(D, device=g.device)
    if path.endswith("/bias"):
        n = {f"{MLP_}/fc1/bias": F, "lm_head/bias": V}.get(path, D)
        return torch.zeros(n, device=g.device)
    if path.endswith("fc1/kernel"):
        return _lecun_normal((D, F), D, g)
    if path.endswith("fc2/kernel"):
        return _lecun_normal((F, D), F, g)
    if path == "lm_head/kernel":
        return _lecun_normal((D, V), D, g)
    if path.endswith("/kernel"):
        return _lecun_normal((D, D), D, g)
    raise KeyError(path)


def init_params(store: ParamStore, seed: int, pp_stage_keys: bool = False, stage_index: int = 0) -> None:
    """Fill the master buffer.  Every logical leaf is generated in full from a key that depends only on
    (seed, path, layer) and then sliced, so DP, TP and PP("shared") start from identical weights.
    ``pp_stage_keys`` reproduces the reference's PP behaviour of drawing stage weights from per-stage keys
    (train/train.py:120-121,144 — quirk Q6): the key then depends on (stage, local layer)."""
    lo, cfg = store.local, store.local.cfg
    r, n = lo.tp_rank, lo.tp_size
    Dl, Fl, Vl = lo.d_local, lo.ff_local, lo.vocab_local
    p = store.p

    dev = store.device

    def leaf(path, layer=None):
        if pp_stage_keys and layer is not None:
            return init_logical_leaf(cfg, path, layer - lo.layer_start, seed, ("pp-stage", stage_index), dev)
        if pp_stage_keys:
            return init_logical_leaf(cfg, path, None, seed, ("pp",), dev)
        return init_logical_leaf(cfg, path, layer, seed, (), dev)

    def put(dst: torch.Tensor, src: torch.Tensor) -> None:
        dst.copy_(src.to(dst.device))

    if lo.has_embed:
        put(p["wte"], leaf("wte/embedding"))
        put(p["wpe"], leaf("wpe/embedding"))
    for i, l in enumerate(lo.layers):
        put(p["ln1_g"][i], leaf(f"{BLOCK}/LayerNorm_0/scale", l))
        put(p["ln2_g"][i], leaf(f"{BLOCK}/LayerNorm_1/scale", l))
        for j, proj in enumerate(("q_proj", "k_proj", "v_proj")):
            w = leaf(f"{ATTN}/{proj}/kernel", l)
            put(p["wqkv"][i][:, j * Dl:(j + 1) * Dl], w[:, r * Dl:(r + 1) * Dl])
        put(p["wo"][i], leaf(f"{ATTN}/out_proj/kernel", l)[r * Dl:(r + 1) * Dl, :])
        put(p["w1"][i], leaf(f"{MLP_}/fc1/kernel", l)[:, r * Fl:(r + 1) * Fl])
        put(p["w2"][i], leaf(f"{MLP_}/fc2/kernel", l)[r * Fl:(r + 1) * Fl, :])
        # biases and LayerNorm biases are zero-initialised: the zero-filled buffers are already right
    if lo.has_head:
        put(p["lnf_g"], leaf("LayerNorm_0/scale"))
        w = leaf("lm_head/kernel")
        v0, nv = lo.vocab_start, lo.vocab_valid
        if nv > 0:
            put(p["wlm"][:, :nv], w[:, v0:v0 + nv])
    store.sync_compute_copy()


# ----------------------------------------------------------------------------------------------------------
# Flax-layout export / import (checkpoint compatibility)
# ----------------------------------------------------------------------------------------------------------
def export_flax_tree(parts: list[tuple[LocalShape, dict[str, torch.Tensor]]]) -> dict[str, np.ndarray]:
    """Assemble the reference-layout tree from per-rank physical tensors.

    ``parts``: one ``(LocalShape, {physical name: tensor})`` per (pipeline stage, tensor-parallel rank).
    Returns path → fp32 ndarray with full, unsharded, UNPADDED shapes and layers stacked on axis 0."""
    cfg = parts[0][0].cfg
    tp = parts[0][0].tp_size
    stages: dict[int, dict[int, tuple[LocalShape, dict[str, torch.Tensor]]]] = {}
    for lo, t in parts:
        stages.setdefault(lo.layer_start, {})[lo.tp_rank] = (lo, t)
    out: dict[str, np.ndarray] = {}

    def cat_tp(group, name, dim):
        return torch.cat([group[r][1][name].detach().float().cpu() for r in range(tp)], dim=dim)

    def rep(group, name):
        return group[0][1][name].detach().float().cpu()

    lay: dict[str, list[torch.Tensor]] = {}
    for start in sorted(stages):
        group = stages[start]
        lo = group[0][0]
        Dl = lo.d_local
        if lo.has_embed:
            out["wte/embedding"] = rep(group, "wte").numpy()
            out["wpe/embedding"] = rep(group, "wpe").numpy()
        if lo.L > 0:
            lay.setdefault(f"{BLOCK}/LayerNorm_0/scale", []).append(rep(group, "ln1_g"))
            lay.setdefault(f"{BLOCK}/LayerNorm_0/bias", []).append(rep(group, "ln1_b"))
            lay.setdefault(f"{BLOCK}/LayerNorm_1/scale", []).append(rep(group, "ln2_g"))
            lay.setdefault(f"{BLOCK}/LayerNorm_1/bias", []).append(rep(group, "ln2_b"))
            for j, proj in enumerate(("q_proj", "k_proj", "v_proj")):
                w = torch.cat([group[r][1]["wqkv"].detach().float().cpu()[:, :, j * Dl:(j + 1) * Dl] for r in range(tp)], 2)
                b = torch.cat([group[r][1]["bqkv"].detach().float().cpu()[:, j * Dl:(j + 1) * Dl] for r in range(tp)], 1)
                lay.setdefault(f"{ATTN}/{proj}/kernel", []).append(w)
                lay.setdefault(f"{ATTN}/{proj}/bias", []).append(b)
            lay.setdefault(f"{ATTN}/out_proj/kernel", []).append(cat_tp(group, "wo", 1))
            lay.setdefault(f"{ATTN}/out_proj/bias", []).append(rep(group, "bo"))
            lay.setdefault(f"{MLP_}/fc1/kernel", []).append(cat_tp(group, "w1", 2))
            lay.setdefault(f"{MLP_}/fc1/bias", []).append(cat_tp(group, "b1", 1))
            lay.setdefault(f"{MLP_}/fc2/kernel", []).append(cat_tp(group, "w2", 1))
            lay.setdefault(f"{MLP_}/fc2/bias", []).append(rep(group, "b2"))
        if lo.has_head:
            out["LayerNorm_0/scale"] = rep(group, "lnf_g").numpy()
            out["LayerNorm_0/bias"] = rep(group, "lnf_b").numpy()
            out["lm_head/kernel"] = cat_tp(group, "wlm", 1)[:, :cfg.vocab_size].contiguous().numpy()
            out["lm_head/bias"] = cat_tp(group, "blm", 0)[:cfg.vocab_size].contiguous().numpy()
    for path, pieces in lay.items():
        out[path] = torch.cat(pieces, 0).numpy()
    return out


def import_flax_tree(tree: dict[str, np.ndarray], local: LocalShape, views: dict[str, torch.Tensor]) -> None:
    """Inverse of :func:`export_flax_tree` for one rank: slice / fuse / pad the logical tree into ``views``."""
    r, Dl, Fl = local.tp_rank, local.d_local, local.ff_local
    ls = slice(local.layers.start, local.layers.stop)

    def t(path):
        return torch.from_numpy(np.asarray(tree[path], dtype=np.float32))

    def put(name, src):
        views[name].copy_(src.to(views[name].device))

    if local.has_embed:
        put("wte", t("wte/embedding"))
        put("wpe", t("wpe/embedding"))
    if local.L > 0:
        put("ln1_g", t(f"{BLOCK}/LayerNorm_0/scale")[ls])
        put("ln1_b", t(f"{BLOCK}/LayerNorm_0/bias")[ls])
        put("ln2_g", t(f"{BLOCK}/LayerNorm_1/scale")[ls])
        put("ln2_b", t(f"{BLOCK}/LayerNorm_1/bias")[ls])
        w = torch.cat([t(f"{ATTN}/{p}/kernel")[ls][:, :, r * Dl:(r + 1) * Dl] for p in ("q_proj", "k_proj", "v_proj")], 2)
        b = torch.cat([t(f"{ATTN}/{p}/bias")[ls][:, r * Dl:(r + 1) * Dl] for p in ("q_proj", "k_proj", "v_proj")], 1)
        put("wqkv", w)
        put("bqkv", b)
        put("wo", t(f"{ATTN}/out_proj/kernel")[ls][:, r * Dl:(r + 1) * Dl, :])
        put("bo", t(f"{ATTN}/out_proj/bias")[ls])
        put("w1", t(f"{MLP_}/fc1/kernel")[ls][:, :, r * Fl:(r + 1) * Fl])
        put("b1", t(f"{MLP_}/fc1/bias")[ls][:, r * Fl:(r + 1) * Fl])
        put("w2", t(f"{MLP_}/fc2/kernel")[ls][:, r * Fl:(r + 1) * Fl, :])
        put("b2", t(f"{MLP_}/fc2/bias")[ls])
    if local.has_head:
        put("lnf_g", t("LayerNorm_0/scale"))
        put("lnf_b", t("LayerNorm_0/bias"))
        v0, nv = local.vocab_start, local.vocab_valid
        views["wlm"].zero_()
        views["blm"].zero_()
        if nv > 0:
            views["wlm"][:, :nv].copy_(t("lm_head/kernel")[:, v0:v0 + nv])
            views["blm"][:nv].copy_(t("lm_head/bias")[v0:v0 + nv])


# ----------------------------------------------------------------------------------------------------------
# The reference's PIPELINE state layout (train/train.py:141-167)
# ----------------------------------------------------------------------------------------------------------
def export_pp_tree(tree: dict[str, np.ndarray], n_stages: int, layer_ranges: list[range] | None = None,
                   dead_copy_scale: float = 1.0) -> dict[str, np.ndarray] | list[dict[str, np.ndarray]]:
    """Re-arrange a merged DP/TP-layout tree (``export_flax_tree``) into the state the reference's ``train_pp`` holds.

    The reference builds, for every stage ``i``, ``{"embed": {wte, wpe}, "stage": {ScanScanBlock_0/… with a leading axis of
    layers_per_stage}, "head": {LayerNorm_0, lm_head}}`` (train/train.py:155-161) and stacks the S per-stage trees on a
    new leading axis with ``jax.device_put_sharded`` (train/train.py:166-167).  Keys here are the ``/``-joined paths of
    that tree (``embed/wte/embedding``, ``stage/ScanScanBlock_0/…``, ``head/lm_head/kernel``).

    ``layer_ranges``: layers owned by each stage; default = the reference's ``n_layers // n_stages`` contiguous layers per
    stage (remainder dropped, quirk Q2).  When every stage owns the same number of layers the result is ONE dict whose
    leaves carry the leading ``S`` axis (the reference's layout); otherwise — this framework's uneven split — a list of
    per-stage dicts, which cannot be stacked.

    ``dead_copy_scale``: the reference keeps a copy of embed / head on EVERY stage although only the first / last stage
    uses them; the unused copies receive zero gradients and merely shrink by weight decay, ``p·(1 − lr·wd)`` per step
    (SURVEY.md §2.7 "PP dead copies").  This framework stores embed / head on their owner stage only; pass
    ``(1 − lr·wd) ** steps`` to reproduce the dead copies' values analytically relative to the INITIAL weights — or keep
    1.0 to fill every copy with the owner's current value."""
    n_layers = next(v.shape[0] for k, v in tree.items() if k.startswith(BLOCK + "/"))
    if layer_ranges is None:
        per = n_layers // n_stages
        layer_ranges = [range(s * per, (s + 1) * per) for s in range(n_stages)]
    if len(layer_ranges) != n_stages:
        raise ValueError(f"{len(layer_ranges)} layer ranges for {n_stages} stages")
    embed_keys = ("wte/embedding", "wpe/embedding")
    head_keys = ("LayerNorm_0/scale", "LayerNorm_0/bias", "lm_head/kernel", "lm_head/bias")
    per_stage: list[dict[str, np.ndarray]] = []
    for s, rng in enumerate(layer_ranges):
        d: dict[str, np.ndarray] = {}
        for k in embed_keys:
            d[f"embed/{k}"] = tree[k] if s == 0 else tree[k] * np.float32(dead_copy_scale)
        for k, v in tree.items():
            if k.startswith(BLOCK + "/"):
                d[f"stage/{k}"] = v[rng.start:rng.stop]
        for k in head_keys:
            d[f"head/{k}"] = tree[k] if s == n_stages - 1 else tree[k] * np.float32(dead_copy_scale)
        per_stage.append(d)
    if len({len(r) for r in layer_ranges}) == 1:
        return {k: np.stack([d[k] for d in per_stage], 0) for k in per_stage[0]}
    return per_stage


def import_pp_tree(pp_tree: dict[str, np.ndarray] | list[dict[str, np.ndarray]]) -> dict[str, np.ndarray]:
    """Inverse of :func:`export_pp_tree`: merge a reference-layout pipeline state back into the DP/TP-layout tree
    (embed from stage 0, head from the last stage, layers concatenated in stage order)."""
    if isinstance(pp_tree, dict):
        S = next(iter(pp_tree.values())).shape[0]
        stages = [{k: v[s] for k, v in pp_tree.items()} for s in range(S)]
    else:
        stages = list(pp_tree)
    out: dict[str, np.ndarray] = {}
    for k, v in stages[0].items():
        if k.startswith("embed/"):
            out[k[len("embed/"):]] = v
    for k, v in stages[-1].items():
        if k.startswith("head/"):
            out[k[len("head/"):]] = v
    for k in stages[0]:
        if k.startswith("stage/"):
            out[k[len("stage/"):]] = np.concatenate([d[k] for d in stages], 0)
    return out
