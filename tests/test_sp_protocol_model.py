"""Protocol-level stress test (randomised delays, repeated-iteration flag reuse) of the folded-barrier
synchronisation of the sequence-parallel TP path (csrc/spsync.cuh, parallel/tp_sp.py with DTC_SP_FOLD=1), modelled with threads.

Every "rank" runs the same kernel sequence as the engine:   LN(push, signal) · wait · [GEMMs] · scatter(signal) ·
LN(wait, reduce, push, signal) · wait · …   with monotonically increasing sequence numbers (k-th signal pairs with the k-th
wait).  There is NO separate "buffer free" handshake: the model checks the claim that the read-after-write flags alone also
prevent the write-after-read hazards —
  * when a LayerNorm step reduces its partial-sum slots, every slot holds exactly the version its producers wrote for THIS step
    (not an older one: RAW; not a newer one: a peer overwrote it too early, WAR);
  * when a GEMM step reads the gathered buffer, every row block holds this step's version.
Random sleeps make ranks overtake each other as far as the protocol allows."""
import random
import threading
import time

import pytest


class Rank:
    def __init__(self, r, n):
        self.r, self.n = r, n
        self.sig = [0] * n            # sig[src]: last sequence number published by rank src (written remotely)
        self.signals = 0              # counters[0]
        self.waits = 0                # counters[1]
        self.part = [None] * n        # partial-sum slot written by rank src
        self.gathered = [None] * n    # row block pushed by rank src


def run_rank(me: Rank, ranks, steps: int, rng: random.Random, errors: list, max_lead: list, step_barrier,
             slow_tail: bool = False):
    n = me.n

    def nap():
        if rng.random() < 0.3:
            time.sleep(rng.random() * 0.002)

    def signal():                     # last CTA: publish the next sequence number to every rank
        me.signals += 1
        for p in ranks:
            p.sig[me.r] = me.signals

    def wait():                       # spin until every rank has published the signal this wait consumes
        me.waits += 1
        deadline = time.time() + 20
        while any(me.sig[s] < me.waits for s in range(n)):
            if time.time() > deadline:
                errors.append(f"rank {me.r}: wait {me.waits} timed out with flags {me.sig}")
                raise TimeoutError
            time.sleep(0)
        max_lead[0] = max(max_lead[0], max(me.sig) - me.waits)

    try:
        version = 0
        for _ in range(steps):
            # the ONE real barrier of a step ("every rank has left the previous step", SeqParallelEngine.forward): the first
            # LayerNorm pushes without waiting for anything, so without it a fast rank would overwrite buffers that a slow
            # rank still reads at the end of its backward pass
            if step_barrier is not None:
                step_barrier.wait(timeout=60)
            # first LayerNorm of a step: nothing to reduce, push my rows, signal
            version += 1
            for p in ranks:
                p.gathered[me.r] = version
            signal()
            for _ in range(4):        # sublayer boundaries of the step
                nap()
                wait()                # (wait kernel) in front of the GEMMs that read the gathered buffer
                if any(g != version for g in me.gathered):
                    errors.append(f"rank {me.r}: gathered buffer {me.gathered} != version {version}")
                nap()
                version += 1
                for p in ranks:       # scatter GEMM: my partial tile into every owner's slot, then signal
                    p.part[me.r] = version
                signal()
                nap()
                wait()                # LayerNorm prologue: all partial slots of THIS step have landed
                if any(v != version for v in me.part):
                    errors.append(f"rank {me.r}: partial slots {me.part} != version {version} (RAW or WAR violation)")
                nap()
                version += 1
                for p in ranks:       # ... reduce, normalise, push my rows to every rank, signal
                    p.gathered[me.r] = version
                signal()
            wait()                    # consume the last push of the step (embedding backward reads the gathered gradient)
            if slow_tail and me.r == 0:
                time.sleep(0.01)      # rank 0 is late reading its buffers: the others are free to run ahead
            if any(g != version for g in me.gathered):
                errors.append(f"rank {me.r}: final gathered buffer {me.gathered} != version {version}")
    except TimeoutError:
        pass


def _simulate(n: int, steps: int, with_step_barrier: bool, slow_tail: bool = False):
    ranks = [Rank(r, n) for r in range(n)]
    errors, max_lead = [], [0]
    barrier = threading.Barrier(n) if with_step_barrier else None
    threads = [threading.Thread(target=run_rank,
                                args=(ranks[r], ranks, steps, random.Random(1000 * n + r), errors, max_lead, barrier, slow_tail))
               for r in range(n)]
    [t.start() for t in threads]
    [t.join(120) for t in threads]
    assert not any(t.is_alive() for t in threads), "deadlock"
    return ranks, errors, max_lead[0]


@pytest.mark.parametrize("n", [2, 4, 8])
def test_raw_flags_alone_keep_partial_slots_and_gathered_buffers_consistent(n):
    ranks, errors, lead = _simulate(n, 25, with_step_barrier=True, slow_tail=(n == 4))
    assert not errors, errors[:3]
    assert all(r.signals == r.waits for r in ranks)          # every signal consumed
    assert lead <= 1                                          # a producer is never more than one signal ahead of a consumer


def test_the_step_start_barrier_is_needed():
    """Without the barrier at the start of a step the model DOES catch the hazard (a fast rank's first push of step k+1 lands
    in a buffer a slow rank still reads at the end of step k) — which is why SeqParallelEngine.forward keeps that one barrier
    kernel in the folded mode too."""
    _, errors, _ = _simulate(4, 10, with_step_barrier=False, slow_tail=True)
    assert errors and all("final gathered buffer" in e for e in errors)
