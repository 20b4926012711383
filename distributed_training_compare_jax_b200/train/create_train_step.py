"""Step builders (component C11; reference: train/create_train_step.py:15-195).

``create_train_step(parallel)`` builds the DP / TP (and hybrid DP×TP) step: forward + backward of the whole
model on this rank's shard, gradient reduction, clip + AdamW — the explicit counterpart of the reference's
jitted ``value_and_grad(loss_fn)`` + ``state.apply_gradients`` whose collectives GSPMD inserted implicitly.
``create_pp_train_step`` builds the pipeline step: a GPipe or 1F1B schedule over microbatches with NCCL
send/recv between stages, per-stage (or global) clipping and a per-stage AdamW, then the loss of the last
stage is broadcast to every stage (the reference's ``psum``, X12).

Both return ``train_step(state, batch, rng) -> (state, loss)`` like the reference; ``loss`` is a device scalar
(fetch it with ``float(loss)`` — the reference's blocking ``np.asarray(loss)``).  On the native backend the
DP/TP step is captured into ONE CUDA graph after a few eager iterations and replayed afterwards (instead of
XLA's compiled executable): the only per-step host work is two small H2D copies and one graph launch.
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import Any

import torch

from ..ops import native
from ..parallel.pp import PipelineExecutor
from ..utils.timing import nvtx_range


@dataclass(frozen=True)
class Batch:
    """x, y: integer token ids [batch, T] (numpy / torch, host or device)."""
    x: Any
    y: Any


@dataclass
class TrainState:
    """Counterpart of flax ``TrainState``: step counter, parameters (inside ``engine.store``) and optimizer."""
    engine: Any
    optimizer: Any
    mesh: Any
    step: int = 0
    batches_consumed: int = 0   # global batches drawn from the token stream so far (saved in checkpoints)
    resumed: bool = False       # state came from a checkpoint: no warm-up steps, the stream continues

    @property
    def params(self):
        return self.engine.store.p

    @property
    def opt_state(self):
        st = self.engine.store
        return {"count": st.step, "mu": st.views_of(st.flat_m), "nu": st.views_of(st.flat_v)}


def _to_i32(a, device, pinned_cache: dict, key: str) -> torch.Tensor:
    """Stage a host batch in a reusable pinned buffer so the H2D copy is asynchronous."""
    t = torch.as_tensor(a)
    if t.dtype != torch.int32:
        t = t.to(torch.int32)
    if device.type != "cuda" or t.is_cuda:
        return t
    buf = pinned_cache.get(key)
    if buf is None or buf.shape != t.shape:
        buf = torch.empty(t.shape, dtype=torch.int32, pin_memory=True)
        pinned_cache[key] = buf
    evt = pinned_cache.get("evt")
    if evt is not None:
        evt.synchronize()  # the previous step's async H2D copy must have consumed the pinned buffer
    buf.copy_(t)
    return buf


def _mark_copied(device, pinned_cache: dict) -> None:
    if device.type == "cuda":
        evt = pinned_cache.get("evt") or torch.cuda.Event()
        evt.record()
        pinned_cache["evt"] = evt


class _GraphedBody:
    """Runs ``body()`` eagerly ``eager_iters`` times, then captures it into a CUDA graph and replays it."""

    def __init__(self, body, enabled: bool, eager_iters: int = 2):
        self.body, self.enabled, self.eager_left = body, enabled, eager_iters
        self.graph = None
        self.launches_per_replay = 0

    def __call__(self) -> None:
        if not self.enabled:
            self.body()
            return
        if self.graph is not None:
            self.graph.replay()
            native.add_launches(self.launches_per_replay)
            return
        if self.eager_left > 0:
            self.eager_left -= 1
            self.body()
            return
        torch.cuda.synchronize()
        g = torch.cuda.CUDAGraph()
        before = native.launches()
        # The step is captured on a HIGH-priority stream: kernel nodes inherit the priority of the stream they were
        # captured on, the engine's side streams (weight gradients, bucketed DP reduction) keep the default (lowest)
        # priority, so whenever both have thread blocks waiting for an SM the dependent chain goes first.
        import os
        prio = os.environ.get("DTC_GRAPH_PRIO", "1") != "0"
        cap = torch.cuda.Stream(priority=-1) if prio else None
        with torch.cuda.graph(g, stream=cap):
            self.body()
        self.launches_per_replay = native.launches() - before  # captured, not executed
        native.add_launches(-self.launches_per_replay)
        self.graph = g
        g.replay()
        native.add_launches(self.launches_per_replay)


def resync_tp_replicas(state: "TrainState") -> None:
    """Make the tensor-parallel ranks' copies of the replicated leaves bit-identical again (tp-rank 0 wins): master
    weights, Adam moments and the bf16 compute copy of the replicated range.  Host-driven, outside the step graph."""
    mesh, st = state.mesh, state.engine.store
    if mesh.tp.size == 1 or st.rep_numel == 0:
        return
    for flat in (st.flat_master, st.flat_m, st.flat_v, st.flat_compute):
        if flat is not None and flat.numel() >= st.rep_numel:
            mesh.tp.broadcast(flat[:st.rep_numel], 0)


def create_train_step(parallel: str = "dp", cuda_graph: bool = True, tp_resync_every: int = 0) -> Any:
    if parallel == "pp":
        raise ValueError("`parallel=pp` must use `create_pp_train_step`")
    cache: dict = {}
    graphed: dict = {}

    def train_step(state: TrainState, batch: Batch, rng: Any = None):
        eng, opt, mesh = state.engine, state.optimizer, state.mesh
        dev = eng.device
        # DP: the loader already yields this rank's slice (reference: P("data", None)); TP: replicated batch
        x, y = _to_i32(batch.x, dev, cache, "x"), _to_i32(batch.y, dev, cache, "y")
        n = eng.n_micro  # > 1: the local batch is split into independent chains that run concurrently on the GPU
        per = x.shape[0] // n
        for m in range(n):
            eng.set_batch(m, x[m * per:(m + 1) * per], y[m * per:(m + 1) * per])
        _mark_copied(dev, cache)

        def body():
            eng.begin_step()
            if n == 1:
                with nvtx_range("forward"):
                    eng.forward(0)
                with nvtx_range("backward"):
                    eng.backward(0)
            else:
                main = torch.cuda.current_stream()
                if "chains" not in cache:
                    cache["chains"] = [torch.cuda.Stream(device=dev) for _ in range(n)]
                for m, st in enumerate(cache["chains"]):
                    st.wait_stream(main)
                    with torch.cuda.stream(st):
                        eng.forward(m)
                        eng.backward(m)
                for st in cache["chains"]:
                    main.wait_stream(st)
            with nvtx_range("optimizer"):
                opt.step()
            mesh.dp.all_reduce(eng.loss_sum)  # mean over the GLOBAL batch (reference: loss.mean() over P("data"))

        if id(eng) not in graphed:
            graphed[id(eng)] = _GraphedBody(body, enabled=cuda_graph and eng.backend == "native")
        graphed[id(eng)]()
        state.step += 1
        if tp_resync_every and mesh.tp.size > 1 and state.step % tp_resync_every == 0 \
                and not hasattr(opt, "sync_full_state"):
            resync_tp_replicas(state)
        return state, eng.loss_sum[0]

    return train_step


def create_pp_train_step(num_microbatches: int, num_stages: int, schedule: str = "gpipe", cuda_graph: bool = False) -> Any:
    cache: dict = {}
    executors: dict = {}

    def train_step(state: TrainState, batch: Batch, rng: Any = None):
        eng, opt, mesh = state.engine, state.optimizer, state.mesh
        dev = eng.device
        x = _to_i32(batch.x, dev, cache, "x")
        y = _to_i32(batch.y, dev, cache, "y")
        mbs = x.shape[0] // num_microbatches
        # every stage receives the whole batch (reference in_axes=None); only the first / last stage use x / y
        for m in range(num_microbatches):
            eng.set_batch(m, x[m * mbs:(m + 1) * mbs], y[m * mbs:(m + 1) * mbs])
        _mark_copied(dev, cache)
        if id(eng) not in executors:
            executors[id(eng)] = PipelineExecutor(eng, mesh.pp, schedule, cuda_graph=cuda_graph)
        eng.begin_step()
        executors[id(eng)].run()
        opt.step()
        mesh.pp.broadcast(eng.loss_sum, num_stages - 1)
        state.step += 1
        return state, eng.loss_sum[0]

    return train_step
