"""Sequence-parallel TP kernels on ONE GPU: the "ranks" are separate local buffers, so the GEMM scatter epilogue, the
reduce + LayerNorm + push kernels and the whole row-sharded engine data flow are checked against the fp32 PyTorch
restatement without a second device (the cross-rank barriers and the multicast stores are covered by
``test_multigpu.py``)."""
import ctypes as C

import pytest
import torch

from distributed_training_compare_jax_b200.ops import functional as Fn
from distributed_training_compare_jax_b200.ops import reference as R

pytestmark = pytest.mark.gpu
BF16, F32 = torch.bfloat16, torch.float32


@pytest.fixture(autouse=True)
def _seed():
    torch.manual_seed(4321)
    yield
    torch.cuda.synchronize()


def rel_err(got, want):
    got, want = got.float(), want.float()
    return ((got - want).norm() / want.norm().clamp_min(1e-12)).item()


def randn(*shape, scale=1.0):
    return (torch.randn(*shape, device="cuda") * scale).to(BF16)


def ptr_array(tensors):
    return (C.c_ulonglong * len(tensors))(*[t.data_ptr() for t in tensors])


@pytest.mark.parametrize("b_mn", [True, False])
@pytest.mark.parametrize("n,M,N,K,bn", [(4, 1024, 512, 256, 0), (2, 512, 256, 320, 128), (8, 4096, 512, 64, 0),
                                        (1, 256, 512, 128, 64)])
def test_gemm_scatter_epilogue(n, M, N, K, bn, b_mn):
    a = randn(M, K)
    w = randn(K, N, scale=K ** -0.5) if b_mn else randn(N, K, scale=K ** -0.5)
    want = a.float() @ (w.float() if b_mn else w.float().t())
    Ms = M // n
    for rank in sorted({0, n - 1}):
        bufs = [torch.zeros(n, Ms, N, device="cuda", dtype=BF16) for _ in range(n)]   # every "rank"'s partial-sum slots
        Fn.gemm_scatter(a, w, ptr_array(bufs), n, rank, N, b_mn=b_mn, block_n=bn)
        torch.cuda.synchronize()
        for o in range(n):
            assert rel_err(bufs[o][rank], want[o * Ms:(o + 1) * Ms]) < 1e-2, (rank, o)
            others = [s for s in range(n) if s != rank]
            if others:
                assert float(bufs[o][others].abs().max()) == 0.0   # only this rank's slot is written


@pytest.mark.parametrize("n_part,n_out,Ms,D", [(0, 1, 128, 512), (2, 2, 256, 512), (8, 3, 64, 256), (4, 1, 100, 1024)])
def test_sp_ln_fwd(n_part, n_out, Ms, D):
    res, bias = randn(Ms, D, scale=2.0), randn(D, scale=0.5)
    part = randn(max(n_part, 1), Ms, D)
    g, b = (1 + 0.1 * torch.randn(D, device="cuda")).to(BF16), randn(D, scale=0.1)
    row0, Mfull = Ms, 3 * Ms
    outs = [torch.full((Mfull, D), 7.0, device="cuda", dtype=BF16) for _ in range(n_out)]
    hsum = torch.empty(Ms, D, device="cuda", dtype=BF16)
    mean, rstd = torch.empty(Ms, device="cuda"), torch.empty(Ms, device="cuda")
    use_bias = n_part > 0
    Fn.sp_ln_fwd(res, bias if use_bias else None, part if n_part else None, n_part, g, b, hsum if n_part else None, mean,
                 rstd, 0, ptr_array(outs), n_out, row0)
    torch.cuda.synchronize()
    x = res.float() + (bias.float() if use_bias else 0) + (part[:n_part].float().sum(0) if n_part else 0)
    xb = x.to(BF16)
    if n_part:
        assert rel_err(hsum, x) < 4e-3
    wy, wmean, wrstd = R.layernorm_fwd(xb, g, b)
    assert rel_err(mean, wmean) < 1e-3 and rel_err(rstd, wrstd) < 1e-3
    for o in outs:
        assert rel_err(o[row0:row0 + Ms], wy) < 1e-2
        assert float((o[:row0].float() - 7).abs().max()) == 0 and float((o[row0 + Ms:].float() - 7).abs().max()) == 0


@pytest.mark.parametrize("n_part,n_out,Ms,D,with_res", [(1, 1, 128, 512, True), (2, 2, 256, 512, False),
                                                        (8, 2, 72, 256, True), (4, 1, 40, 1024, True)])
def test_sp_ln_bwd(n_part, n_out, Ms, D, with_res):
    x, part = randn(Ms, D, scale=2.0), randn(n_part, Ms, D)
    g = (1 + 0.1 * torch.randn(D, device="cuda")).to(BF16)
    dres = randn(Ms, D) if with_res else None
    _, mean, rstd = R.layernorm_fwd(x, g, torch.zeros_like(g))
    mean, rstd = mean.contiguous(), rstd.contiguous()
    dg, db = torch.zeros(D, device="cuda"), torch.zeros(D, device="cuda")
    row0, Mfull = 2 * Ms, 3 * Ms
    outs = [torch.full((Mfull, D), -3.0, device="cuda", dtype=BF16) for _ in range(n_out)]
    Fn.sp_ln_bwd(part, n_part, x, mean, rstd, g, dres, dg, db, 0, ptr_array(outs), n_out, row0)
    torch.cuda.synchronize()
    wdx, wdg, wdb = R.layernorm_bwd(part.float().sum(0), x, mean, rstd, g, dres=dres)
    assert rel_err(dg, wdg) < 1e-2 and rel_err(db, wdb) < 1e-2
    for o in outs:
        assert rel_err(o[row0:], wdx) < 1e-2
        assert float((o[:row0].float() + 3).abs().max()) == 0


def _engines(dropout):
    from distributed_training_compare_jax_b200.config.schema import ModelConfig
    from distributed_training_compare_jax_b200.models.param_tree import LocalShape, ParamStore, init_params
    from distributed_training_compare_jax_b200.parallel.mesh import single_device_mesh
    from distributed_training_compare_jax_b200.parallel.tp_sp import SeqParallelEngine
    from distributed_training_compare_jax_b200.train.engine import NativeStageEngine

    mc = ModelConfig(vocab_size=5000, d_model=256, n_layers=3, n_heads=8, d_ff=512, max_seq_len=128, dropout=dropout,
                     parallel="tp")
    mesh = single_device_mesh()
    local = LocalShape(mc)
    B, T = 2, mc.max_seq_len
    out = []
    for cls in (NativeStageEngine, SeqParallelEngine):
        store = ParamStore(local, mesh.device)
        init_params(store, 0)
        store.sync_compute_copy()
        out.append(cls(mc, local, mesh, store, B, T, 1, B * T, 0))
    return mc, out[0], out[1]


@pytest.mark.parametrize("dropout", [0.0, 0.1])
def test_seq_parallel_engine_single_rank_matches_native_engine(dropout):
    """tp = 1: every partial-sum slot, gathered buffer and push is local, so the row-sharded forward / backward must
    reproduce the replicated-activation engine (same kernels elsewhere) up to the extra bf16 rounding of the partials."""
    mc, base, sp = _engines(dropout)
    g = torch.Generator().manual_seed(0)
    tok = torch.randint(0, mc.vocab_size - 1, (2, mc.max_seq_len + 1), generator=g, dtype=torch.int32).cuda()
    x, y = tok[:, :-1].contiguous(), tok[:, 1:].contiguous()
    for e in (base, sp):
        e.begin_step()
        e.set_batch(0, x, y)
        e.forward(0)
        e.backward(0)
    torch.cuda.synchronize()
    lb, ls = base.loss_sum.item(), sp.loss_sum.item()
    assert abs(lb - ls) < 5e-3 * abs(lb), (lb, ls)
    for name in base.store.g:
        a, w = sp.store.g[name].float(), base.store.g[name].float()
        err = ((a - w).norm() / w.norm().clamp_min(1e-12)).item()
        assert err < 3e-2, f"{name}: rel err {err:.4f}"
