"""CPU model of the attention forward kernel's block -> work mapping (csrc/attention.cu, flash_attn_fwd_pipe_kernel):
CTAs are dispatched in linear block order and two fit an SM (256 TMEM columns each), so the causal work — query block q of a
(batch, head) pair visits 2 * (q + 1) key blocks — is ordered heaviest-first over the WHOLE grid.  The model checks that the
index arithmetic is a bijection and that the order shortens the makespan of a greedy slot scheduler."""
import heapq
import itertools

import pytest


def mapping(lin, gx, H, B):
    """mirror of the kernel: grid (gx = T / 128, H, B), lin = x + gx * (y + H * z)"""
    n_bh = H * B
    qblk = gx - 1 - lin // n_bh
    h = (lin % n_bh) % H
    b = (lin % n_bh) // H
    return qblk, h, b


def interleaved(lin, gx, H, B):
    """the previous order: heaviest first inside each (batch, head) pair only"""
    x, rest = lin % gx, lin // gx
    return gx - 1 - x, rest % H, rest // H


def makespan(order, slots):
    """greedy: CTAs start in dispatch order as slots free up; cost of a CTA = key blocks it visits"""
    free = [0] * slots
    heapq.heapify(free)
    end = 0
    for qblk in order:
        t = heapq.heappop(free) + 2 * (qblk + 1)
        end = max(end, t)
        heapq.heappush(free, t)
    return end


@pytest.mark.parametrize("gx,H,B", [(4, 16, 8), (4, 16, 1), (1, 16, 1), (8, 16, 2), (2, 3, 5)])
def test_block_mapping_is_a_bijection(gx, H, B):
    seen = {mapping(lin, gx, H, B) for lin in range(gx * H * B)}
    assert seen == set(itertools.product(range(gx), range(H), range(B)))


def test_global_heaviest_first_is_monotone_and_shortens_the_tail():
    gx, H, B, slots = 4, 16, 8, 2 * 148
    order = [mapping(lin, gx, H, B)[0] for lin in range(gx * H * B)]
    assert order == sorted(order, reverse=True)
    old = [interleaved(lin, gx, H, B)[0] for lin in range(gx * H * B)]
    assert sorted(old) == sorted(order)
    assert makespan(order, slots) < makespan(old, slots)
    # 2560 key-block visits on 296 slots: 8.65 per slot at best; heaviest-first finishes within 10
    assert makespan(order, slots) <= 10
