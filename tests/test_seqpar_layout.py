"""Host-side layout of the sequence-parallel TP state (parallel/tp_sp.py) — buffer carving, peer / multicast addressing and the
applicability checks — on CPU tensors (a single-rank group needs no symmetric memory)."""
import types

import pytest
import torch

from distributed_training_compare_jax_b200.config.schema import ModelConfig
from distributed_training_compare_jax_b200.models.param_tree import LocalShape
from distributed_training_compare_jax_b200.parallel.mesh import single_device_mesh
from distributed_training_compare_jax_b200.parallel.tp_sp import SeqParallelState


def _engine(batch=2, T=128, L=3, D=256):
    mc = ModelConfig(vocab_size=1000, d_model=D, n_layers=L, n_heads=8, d_ff=512, max_seq_len=T, dropout=0.0, parallel="tp")
    return types.SimpleNamespace(cfg=mc, local=LocalShape(mc), micro_batch=batch, seq_len=T, device=torch.device("cpu"))


def test_arena_is_carved_into_gathered_operands_gradients_and_partial_slots():
    eng = _engine()
    st = SeqParallelState(eng, single_device_mesh("cpu"))
    M, D, L = 2 * 128, 256, 3
    assert (st.n, st.rank, st.Ms, st.row0) == (1, 0, M, 0)
    assert st.xg1.shape == (L, M, D) and st.xg2.shape == (L, M, D) and st.xgf.shape == (M, D)
    assert st.part.shape == (M, D) and st.arena.tensor.numel() == (2 * L + 5) * M * D
    views = [st.xg1[0], st.xg1[2], st.xg2[1], st.xgf, st.dA, st.dB, st.dMID, st.part]
    ptrs = [v.data_ptr() for v in views]
    assert len(set(ptrs)) == len(ptrs) and all((p - st.arena.tensor.data_ptr()) % (M * D * 2) == 0 for p in ptrs)
    mc_addr, peers, n = st.dest(st.xg2[1])
    assert mc_addr == 0 and n == 1 and list(peers) == [st.xg2[1].data_ptr()]      # no multicast on a single rank
    assert st.dest(st.xg2[1]) is st.dest(st.xg2[1])                                # cached per view
    assert list(st.part_ptrs) == [st.part.data_ptr()]
    assert st.own(st.dA).shape == (st.Ms, D) and st.own(st.dA).data_ptr() == st.dA.data_ptr()
    assert len(st.hs) == L + 1 and len(st.hmids) == L and st.mean1.shape == (L, st.Ms)
    st.barrier()        # no barrier object for a single rank: a no-op
    assert not st.fold


def test_rows_must_split_into_whole_128_row_tiles():
    with pytest.raises(ValueError):
        SeqParallelState(_engine(batch=1, T=64), single_device_mesh("cpu"))      # 64 rows: not a multiple of 128
