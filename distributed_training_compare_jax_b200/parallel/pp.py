"""Pipeline parallelism: schedules + point-to-point executor (components C11-PP / R12; reference:
train/create_train_step.py:55-195 — a GPipe clock loop of ``M + S − 1`` ticks with ``ppermute`` of
(h, y, valid) every tick, and train/train.py:107-233 for the stage layout).

What is kept: one stage per rank, contiguous layers per stage, GPipe (all forwards, then all backwards) as the
default schedule, the loss of the last stage made visible on every stage (reference ``psum``, X12).
What is dropped on purpose: labels and validity flags do not travel with the activations
(every stage gets the batch from its own loader; validity is static in the schedule) and nothing is sent on
bubble ticks.  Added: the 1F1B schedule (at most S − stage microbatches in flight).

Transfers stay on NCCL send/recv (``torch.distributed`` P2P; gloo on CPU).  Activations (downstream) and
activation gradients (upstream) travel on TWO separate communicators, each with its own NCCL stream: a send is
issued right after the action that produced it and is not waited for, a receive is issued right before the
action that needs it.  Because the two directions never queue behind each other, the bidirectional steady
state of 1F1B cannot form a wait cycle, and no batching of unrelated sends and receives is needed (measured on
B200: a batched group holding a send and a receive whose peer issues them as two separate groups never
completes under NCCL 2.28 — ``tools/p2p_probe.py``).
"""
from __future__ import annotations

from dataclasses import dataclass

import torch
import torch.distributed as dist

from ..utils.timing import nvtx_range


@dataclass(frozen=True)
class Action:
    kind: str  # "F" or "B"
    mb: int


def stage_layers(n_layers: int, n_stages: int, reference_truncate: bool = False, head_cost: float = 0.0) -> list[range]:
    """Contiguous layer ranges per stage.  ``reference_truncate`` reproduces the reference's
    ``layers_per_stage = n_layers // n_stages`` (train/train.py:118), which silently drops the remainder
    (12 layers on 8 stages → an 8-layer model, quirk Q2); the default spreads all layers (e.g. 2,2,2,2,1,1,1,1).
    ``head_cost`` > 0 (``pp_layer_split: balanced``): the last stage also runs LayerNorm + lm_head + cross-entropy, which
    cost about ``head_cost`` transformer layers (3.5 for the reference model: 0.95 ms against 0.27 ms per layer); layers are
    then assigned so that the stage COSTS are as equal as contiguous ranges allow."""
    if reference_truncate:
        per = n_layers // n_stages
        return [range(s * per, (s + 1) * per) for s in range(n_stages)]
    if head_cost > 0 and n_stages > 1:
        total = n_layers + head_cost
        counts, used = [], 0
        for s in range(n_stages - 1):
            cum = int(round((s + 1) * total / n_stages))            # layers that should lie at or before this boundary
            n = max(1, min(cum - used, n_layers - used - (n_stages - 1 - s)))   # >= 1 here, >= 1 left for every later stage
            counts.append(n)
            used += n
        counts.append(n_layers - used)
        out, start = [], 0
        for n in counts:
            out.append(range(start, start + n))
            start += n
        return out
    base, extra = divmod(n_layers, n_stages)
    out, start = [], 0
    for s in range(n_stages):
        n = base + (1 if s < extra else 0)
        out.append(range(start, start + n))
        start += n
    return out


def build_schedule(kind: str, n_stages: int, stage: int, n_micro: int) -> list[Action]:
    """Per-stage action list.  gpipe: F0..F(M-1), B(M-1)..B0 (backward in reverse microbatch order, as the
    autodiff of the reference's forward scan does).  1f1b: S-1-stage warm-up forwards, then alternate."""
    if kind == "gpipe":
        return [Action("F", m) for m in range(n_micro)] + [Action("B", m) for m in reversed(range(n_micro))]
    if kind == "1f1b":
        warm = min(n_stages - 1 - stage, n_micro)
        acts = [Action("F", m) for m in range(warm)]
        f, b = warm, 0
        while b < n_micro:
            if f < n_micro:
                acts.append(Action("F", f))
                f += 1
            acts.append(Action("B", b))
            b += 1
        return acts
    raise ValueError(f"unknown pipeline schedule `{kind}`")


def bubble_fraction(n_stages: int, n_micro: int) -> float:
    """(S−1)/(M+S−1): idle share of a GPipe / 1F1B pipeline with equal stage times."""
    return (n_stages - 1) / (n_micro + n_stages - 1)


def max_in_flight(actions: list[Action]) -> int:
    live = peak = 0
    for a in actions:
        live += 1 if a.kind == "F" else -1
        peak = max(peak, live)
    return peak


class P2PBackend:
    """torch.distributed point-to-point over the PP axis (NCCL on GPU, gloo on CPU): one communicator per
    direction (``axis.group`` downstream, ``axis.group2`` upstream)."""

    def __init__(self, axis):
        self.axis = axis
        self.inflight: list = []

    def _group(self, peer: int):
        down = peer > self.axis.rank
        g2 = getattr(self.axis, "group2", None)
        return self.axis.group if (down or g2 is None) else g2

    def send(self, t: torch.Tensor, peer: int) -> None:
        """Asynchronous: the buffer must stay untouched until ``flush`` (end of the step)."""
        self.inflight.append(dist.isend(t, self.axis.ranks[peer], group=self._group_for_send(peer)))

    def _group_for_send(self, peer: int):
        # a downstream send (to stage+1) pairs with the peer's receive from stage-1: both use the downstream group
        return self._group(peer)

    def recv(self, t: torch.Tensor, peer: int) -> None:
        down = peer < self.axis.rank  # data arriving from the previous stage travels downstream
        g2 = getattr(self.axis, "group2", None)
        g = self.axis.group if (down or g2 is None) else g2
        dist.irecv(t, self.axis.ranks[peer], group=g).wait()

    def flush(self) -> None:
        for w in self.inflight:
            w.wait()
        self.inflight = []


class PipelineExecutor:
    """Walks a stage's action list, moving activations (forward) and activation gradients (backward)."""

    def __init__(self, engine, axis, schedule: str, p2p=None, cuda_graph: bool = False, eager_steps: int = 2):
        self.engine, self.axis = engine, axis
        self.stage, self.n_stages = axis.rank, axis.size
        self.actions = build_schedule(schedule, self.n_stages, self.stage, engine.n_micro)
        self.p2p = p2p or P2PBackend(axis)
        self.trace: list[tuple[str, int]] = []  # (event, mb) — used by the schedule tests
        # Optional (native engine): every (forward | backward, microbatch) of this stage becomes one CUDA graph, so a step
        # costs 2·M graph launches + the NCCL send / recv calls instead of a few hundred kernel launches from Python
        # (the host of the GPU box needs ~15 us per launch: an eager pipeline step is host-bound).  All graphs are
        # captured in one pass BETWEEN two steps — capture records without executing and involves no communication, so
        # every stage does it at the same step without any cross-stage ordering constraint.
        self.cuda_graph = bool(cuda_graph and getattr(engine, "backend", "") == "native")
        self.eager_steps = eager_steps
        self.steps_run = 0
        self.graphs: dict[tuple[str, int], tuple] = {}

    def _capture_all(self) -> None:
        from ..ops import native
        eng = self.engine
        torch.cuda.synchronize()
        for kind, mb in sorted({(a.kind, a.mb) for a in self.actions}):
            g = torch.cuda.CUDAGraph()
            before = native.launches()
            with torch.cuda.graph(g):
                out = eng.forward(mb) if kind == "F" else eng.backward(mb)
            n = native.launches() - before          # recorded, not executed
            native.add_launches(-n)
            self.graphs[(kind, mb)] = (g, out, n)

    def _do(self, kind: str, mb: int):
        hit = self.graphs.get((kind, mb))
        if hit is None:
            return self.engine.forward(mb) if kind == "F" else self.engine.backward(mb)
        from ..ops import native
        g, out, n = hit
        g.replay()
        native.add_launches(n)
        return out

    def run(self) -> None:
        eng, first, last = self.engine, self.stage == 0, self.stage == self.n_stages - 1
        if self.cuda_graph and not self.graphs and self.steps_run >= self.eager_steps:
            self._capture_all()
        self.steps_run += 1
        for a in self.actions:
            if a.kind == "F" and not first:
                self.p2p.recv(eng.input_buffer(a.mb), self.stage - 1)
            if a.kind == "B" and not last:
                self.p2p.recv(eng.grad_input_buffer(a.mb), self.stage + 1)
            self.trace.append((a.kind, a.mb))
            if a.kind == "F":
                with nvtx_range(f"F{a.mb}"):
                    out = self._do("F", a.mb)
                if not last:
                    self.p2p.send(out, self.stage + 1)
            else:
                with nvtx_range(f"B{a.mb}"):
                    gin = self._do("B", a.mb)
                if not first:
                    self.p2p.send(gin.contiguous(), self.stage - 1)
        self.p2p.flush()
