"""Checkpoint / resume in the reference's Flax/optax layout.

The reference never serialises anything; BASELINE.json nevertheless asks for a parameter tree compatible with
its Flax/optax state.  A checkpoint is one ``.npz`` whose keys are ``/``-joined Flax paths:

    params/<path>            full, unsharded, unpadded fp32 leaves (kernels (in,out), layers stacked on axis 0)
    opt_state/mu/<path>      Adam first moment  (optax ScaleByAdamState.mu)   — same tree as params
    opt_state/nu/<path>      Adam second moment (optax ScaleByAdamState.nu)
    opt_state/count, step    optimizer step count / TrainState.step
    meta/*                   strategy, mesh_shape (dp, tp, pp), batches_consumed (position in the token stream)

Saving gathers every rank's physical tensors to rank 0 (``gather_object``) and merges them with
``export_flax_tree``; loading slices the full tree for the local rank (``import_flax_tree``), so a checkpoint
written under one strategy / mesh can be resumed under any other."""
from __future__ import annotations

import os

import numpy as np
import torch
import torch.distributed as dist

from ..models.param_tree import ParamStore, export_flax_tree, import_flax_tree


def _gather_parts(store: ParamStore, flat: torch.Tensor, mesh):
    views = {k: v.detach().cpu() for k, v in store.views_of(flat).items()}
    part = (store.local, views)
    if mesh.world_size == 1:
        return [part]
    if mesh.dp.rank != 0:           # data-parallel replicas hold identical state: only dp-rank 0 contributes
        part = None
    gathered = [None] * mesh.world_size if mesh.rank == 0 else None
    dist.gather_object(part, gathered, dst=0)
    return [p for p in gathered if p is not None] if mesh.rank == 0 else None


def save_checkpoint(path: str, store: ParamStore, mesh, step: int, meta: dict | None = None) -> None:
    blobs = {}
    for prefix, flat in (("params", store.flat_master), ("opt_state/mu", store.flat_m), ("opt_state/nu", store.flat_v)):
        if flat is None:
            continue
        parts = _gather_parts(store, flat, mesh)
        if mesh.rank == 0:
            for k, v in export_flax_tree(parts).items():
                blobs[f"{prefix}/{k}"] = v
    if mesh.rank == 0:
        blobs["opt_state/count"] = np.asarray(int(store.step.item()), dtype=np.int32)
        blobs["step"] = np.asarray(step, dtype=np.int32)
        for k, v in (meta or {}).items():
            blobs[f"meta/{k}"] = np.asarray(v)
        os.makedirs(os.path.dirname(os.path.abspath(path)), exist_ok=True)
        tmp = path + ".tmp.npz"
        np.savez(tmp, **blobs)
        os.replace(tmp, path)
    mesh.barrier()


def load_checkpoint(path: str, store: ParamStore) -> dict:
    """Fills the local store (master, Adam moments, step) from a Flax-layout checkpoint; returns the meta dict."""
    z = np.load(path)

    def sub(prefix):
        n = len(prefix) + 1
        return {k[n:]: z[k] for k in z.files if k.startswith(prefix + "/")}

    with torch.no_grad():   # the torch engine has already marked the parameter views as autograd leaves
        import_flax_tree(sub("params"), store.local, store.p)
        if store.flat_m is not None and any(k.startswith("opt_state/mu/") for k in z.files):
            import_flax_tree(sub("opt_state/mu"), store.local, store.views_of(store.flat_m))
            import_flax_tree(sub("opt_state/nu"), store.local, store.views_of(store.flat_v))
    if "opt_state/count" in z.files:
        store.step.fill_(int(z["opt_state/count"]))
    store.sync_compute_copy()
    meta = {k[5:]: z[k] for k in z.files if k.startswith("meta/")}
    meta["step"] = int(z["step"]) if "step" in z.files else 0
    return meta
