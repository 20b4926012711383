"""Grouped per-layer gradient kernels (validated on B200 in round 2; the grouped backward is the engine's default):

  * ``dtc_gemm_wgrad_grouped`` / ``dtc_gemm_wgrad_grouped_pair`` / ``dtc_colsum_grouped`` — the four weight / bias gradients
    of a layer as one launch each (128 x 128 tiles on single CTAs, or 256 x 256 tiles on CTA pairs)
  * ``NativeStageEngine(wgrad_grouped=True)`` against the per-GEMM backward (``wgrad_grouped=False``)
"""
import os

import pytest
import torch

from distributed_training_compare_jax_b200.ops import functional as Fn

pytestmark = pytest.mark.gpu
BF16, F32 = torch.bfloat16, torch.float32


def rel_err(got, want):
    got, want = got.float(), want.float()
    return ((got - want).norm() / want.norm().clamp_min(1e-12)).item()


def randn(*shape, scale=1.0):
    return (torch.randn(*shape, device="cuda") * scale).to(BF16)


@pytest.mark.parametrize("pair", [False, True])
@pytest.mark.parametrize("splits", [1, 3])
@pytest.mark.parametrize("K,shapes", [(4096, [(2048, 512), (512, 2048), (512, 512), (512, 1536)]),
                                      (512, [(256, 768), (200, 264)]), (1000, [(128, 128)])])
def test_wgrad_grouped_matches_fp32_reference(K, shapes, splits, pair):
    torch.manual_seed(7)
    xs = [randn(K, m) for m, _ in shapes]
    dys = [randn(K, n, scale=K ** -0.5) for _, n in shapes]
    outs = [torch.full((m, n), 0.25, device="cuda", dtype=F32) for m, n in shapes]     # accumulated into
    Fn.wgrad_grouped(Fn.wgrad_group_args(xs, dys, outs), splits=splits, pair=pair)
    torch.cuda.synchronize()
    for x, dy, o in zip(xs, dys, outs):
        want = x.float().t() @ dy.float() + 0.25
        assert rel_err(o, want) < 1e-2


def test_colsum_grouped_matches_reference():
    torch.manual_seed(8)
    shapes = [(4096, 512), (4096, 2048), (1000, 264), (64, 1536)]
    xs = [randn(m, n) for m, n in shapes]
    outs = [torch.ones(n, device="cuda", dtype=F32) for _, n in shapes]
    Fn.colsum_grouped(Fn.colsum_group_args(xs, outs))
    torch.cuda.synchronize()
    for x, o in zip(xs, outs):
        assert rel_err(o, x.float().sum(0) + 1) < 1e-2


@pytest.mark.parametrize("L", [1, 2, 5])
def test_grouped_backward_matches_default_backward(L):
    from distributed_training_compare_jax_b200.config.schema import ModelConfig
    from distributed_training_compare_jax_b200.models.param_tree import LocalShape, ParamStore, init_params
    from distributed_training_compare_jax_b200.parallel.mesh import single_device_mesh
    from distributed_training_compare_jax_b200.train.engine import NativeStageEngine

    mc = ModelConfig(vocab_size=5000, d_model=256, n_layers=L, n_heads=8, d_ff=512, max_seq_len=128, dropout=0.1,
                     parallel="dp")
    mesh, local = single_device_mesh(), LocalShape(mc)
    B, T = 2, mc.max_seq_len
    engines = []
    for grouped in (False, True):
        store = ParamStore(local, mesh.device)
        init_params(store, 0)
        store.sync_compute_copy()
        engines.append(NativeStageEngine(mc, local, mesh, store, B, T, 1, B * T, 0, wgrad_grouped=grouped))
    g = torch.Generator().manual_seed(0)
    tok = torch.randint(0, mc.vocab_size - 1, (B, T + 1), generator=g, dtype=torch.int32).cuda()
    for e in engines:
        for _ in range(2):   # twice: the second pass reuses the cached argument blocks and accumulates
            e.begin_step()
            e.set_batch(0, tok[:, :-1].contiguous(), tok[:, 1:].contiguous())
            e.forward(0)
            e.backward(0)
    torch.cuda.synchronize()
    base, grp = engines
    assert abs(base.loss_sum.item() - grp.loss_sum.item()) < 1e-4 * abs(base.loss_sum.item())
    for name in base.store.g:
        err = rel_err(grp.store.g[name], base.store.g[name])
        assert err < 5e-3, f"{name}: rel err {err:.4f}"


def test_sequence_parallel_engine_with_folded_barriers_single_rank():
    """tp = 1, DTC_SP_FOLD mode: every signal / wait pairs up on the one rank, so a counting error in the protocol shows
    up as a (bounded) spin timeout or as wrong gradients; two passes check that the sequence numbers keep advancing."""
    from distributed_training_compare_jax_b200.config.schema import ModelConfig
    from distributed_training_compare_jax_b200.models.param_tree import LocalShape, ParamStore, init_params
    from distributed_training_compare_jax_b200.parallel.mesh import single_device_mesh
    from distributed_training_compare_jax_b200.parallel.tp_sp import SeqParallelEngine

    mc = ModelConfig(vocab_size=5000, d_model=256, n_layers=3, n_heads=8, d_ff=512, max_seq_len=128, dropout=0.1,
                     parallel="tp")
    mesh, local = single_device_mesh(), LocalShape(mc)
    B, T = 2, mc.max_seq_len
    engines = []
    for fold in (False, True):
        store = ParamStore(local, mesh.device)
        init_params(store, 0)
        store.sync_compute_copy()
        engines.append(SeqParallelEngine(mc, local, mesh, store, B, T, 1, B * T, 0, fold_barriers=fold))
    g = torch.Generator().manual_seed(0)
    tok = torch.randint(0, mc.vocab_size - 1, (B, T + 1), generator=g, dtype=torch.int32).cuda()
    for e in engines:
        for _ in range(2):
            e.begin_step()
            e.set_batch(0, tok[:, :-1].contiguous(), tok[:, 1:].contiguous())
            e.forward(0)
            e.backward(0)
    torch.cuda.synchronize()
    plain, fold = engines
    signals, waits, t0, t1 = fold.sp.counters.tolist()
    assert signals == waits and signals > 0 and t0 == 0 and t1 == 0          # every signal consumed, tickets reset
    assert abs(plain.loss_sum.item() - fold.loss_sum.item()) < 1e-4 * abs(plain.loss_sum.item())
    for name in plain.store.g:
        err = rel_err(fold.store.g[name], plain.store.g[name])
        assert err < 5e-3, f"{name}: rel err {err:.4f}"
