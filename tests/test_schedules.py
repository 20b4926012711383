"""T6: pipeline schedules (reference: the M+S-1 tick GPipe loop of train/create_train_step.py:170-182) —
event order, in-flight bounds, bubble arithmetic and deadlock freedom with a fake in-process P2P backend."""
import threading

import pytest
import torch

from distributed_training_compare_jax_b200.parallel.mesh import Axis
from distributed_training_compare_jax_b200.parallel.pp import (Action, PipelineExecutor, bubble_fraction,
                                                              build_schedule, max_in_flight, stage_layers)


def test_stage_layers_even_uneven_and_reference_truncation():
    assert [list(r) for r in stage_layers(12, 2)] == [list(range(6)), list(range(6, 12))]
    assert [len(r) for r in stage_layers(12, 8)] == [2, 2, 2, 2, 1, 1, 1, 1]          # all 12 layers kept
    assert sum(len(r) for r in stage_layers(12, 8, reference_truncate=True)) == 8      # reference quirk Q2
    assert [len(r) for r in stage_layers(12, 8, reference_truncate=True)] == [1] * 8


@pytest.mark.parametrize("S,M", [(2, 2), (4, 8), (8, 2), (3, 5)])
def test_gpipe_and_1f1b_orders(S, M):
    for stage in range(S):
        g = build_schedule("gpipe", S, stage, M)
        assert g == [Action("F", m) for m in range(M)] + [Action("B", m) for m in reversed(range(M))]
        assert max_in_flight(g) == M
        o = build_schedule("1f1b", S, stage, M)
        assert sorted(a.mb for a in o if a.kind == "F") == list(range(M))
        assert [a.mb for a in o if a.kind == "B"] == list(range(M))
        seen = set()
        for a in o:                       # a backward only after its own forward
            if a.kind == "F":
                seen.add(a.mb)
            else:
                assert a.mb in seen
        assert max_in_flight(o) <= min(S - stage, M)
    with pytest.raises(ValueError):
        build_schedule("zigzag", S, 0, M)


def test_bubble_fraction_matches_survey_numbers():
    assert abs(bubble_fraction(2, 2) - 1 / 3) < 1e-12      # S=2, M=2 -> 33 %
    assert abs(bubble_fraction(8, 2) - 7 / 9) < 1e-12      # S=8, M=2 -> 78 %


class _FakeNet:
    """Rendezvous mailboxes between threads: a send blocks until the matching receive has been posted."""

    def __init__(self):
        self.lock = threading.Condition()
        self.box = {}

    def send(self, me, t, peer):
        with self.lock:
            self.box.setdefault((me, peer), []).append(t.clone())
            self.lock.notify_all()

    def recv(self, me, t, peer):
        with self.lock:
            ok = self.lock.wait_for(lambda: self.box.get((peer, me)), timeout=5)
            assert ok, f"stage {me} starved waiting for stage {peer}: deadlock"
            t.copy_(self.box[(peer, me)].pop(0))


class _FakeP2P:
    def __init__(self, net, me):
        self.net, self.me = net, me

    def send(self, t, peer):
        self.net.send(self.me, t, peer)

    def recv(self, t, peer):
        self.net.recv(self.me, t, peer)

    def flush(self):
        pass


class _FakeEngine:
    """Forward adds 1 per stage, backward adds 10 per stage — lets the test check what travelled where."""

    def __init__(self, stage, S, M):
        self.stage, self.S, self.n_micro = stage, S, M
        self.inp = [torch.zeros(2) for _ in range(M)]
        self.gin = [torch.zeros(2) for _ in range(M)]
        self.final = {}

    def input_buffer(self, mb):
        return self.inp[mb]

    def grad_input_buffer(self, mb):
        return self.gin[mb]

    def forward(self, mb):
        x = (torch.full((2,), float(mb)) if self.stage == 0 else self.inp[mb]) + 1
        if self.stage == self.S - 1:
            self.final[mb] = x.clone()
            return None
        return x

    def backward(self, mb):
        g = (torch.zeros(2) if self.stage == self.S - 1 else self.gin[mb]) + 10
        if self.stage == 0:
            self.final[mb] = g.clone()
            return None
        return g


@pytest.mark.parametrize("kind", ["gpipe", "1f1b"])
@pytest.mark.parametrize("S,M", [(2, 2), (4, 6), (3, 1)])
def test_executor_moves_data_without_deadlock(kind, S, M):
    net = _FakeNet()
    engines = [_FakeEngine(s, S, M) for s in range(S)]
    execs = [PipelineExecutor(engines[s], Axis("pp", S, s, None, list(range(S))), kind, p2p=_FakeP2P(net, s))
             for s in range(S)]
    errs = []

    def run(e):
        try:
            e.run()
        except BaseException as ex:  # noqa: BLE001
            errs.append(ex)

    ts = [threading.Thread(target=run, args=(e,)) for e in execs]
    [t.start() for t in ts]
    [t.join(30) for t in ts]
    assert not errs, errs
    for mb in range(M):
        assert torch.all(engines[-1].final[mb] == mb + S)        # went through S forwards
        assert torch.all(engines[0].final[mb] == 10 * S)          # and S backwards
    for e in execs:
        assert len(e.trace) == 2 * M


def test_per_action_cuda_graphs_are_only_used_with_the_native_engine():
    # the option is a no-op for the torch / fake engines: the executor keeps calling forward / backward directly
    net = _FakeNet()
    engines = [_FakeEngine(s, 2, 2) for s in range(2)]
    execs = [PipelineExecutor(engines[s], Axis("pp", 2, s, None, [0, 1]), "gpipe", p2p=_FakeP2P(net, s), cuda_graph=True)
             for s in range(2)]
    assert not any(e.cuda_graph for e in execs)
    for _ in range(4):      # past the eager warm-up steps: still no capture
        ts = [threading.Thread(target=e.run) for e in execs]
        [t.start() for t in ts]
        [t.join(30) for t in ts]
    assert all(not e.graphs and e.steps_run == 4 for e in execs)
    assert torch.equal(engines[1].final[0], torch.full((2,), 2.0))


def test_balanced_layer_split_moves_layers_off_the_head_stage():
    """pp_layer_split: balanced — the last stage also runs lm_head + CE (~3.5 layers of work for the reference model)."""
    from distributed_training_compare_jax_b200.parallel.pp import stage_layers
    assert [len(r) for r in stage_layers(12, 2)] == [6, 6]                      # reference: equal layer counts
    assert [len(r) for r in stage_layers(12, 2, False, 3.5)] == [8, 4]          # 8 vs 4 + 3.5
    for S in (2, 3, 4, 8, 12):
        rs = stage_layers(12, S, False, 3.5)
        assert sum(len(r) for r in rs) == 12 and all(len(r) >= 1 for r in rs)
        assert [r.start for r in rs] == [sum(len(q) for q in rs[:i]) for i in range(S)]   # contiguous, in order
        even = stage_layers(12, S)
        cost = lambda split: max(len(r) + (3.5 if i == S - 1 else 0) for i, r in enumerate(split))
        assert cost(rs) <= cost(even)
