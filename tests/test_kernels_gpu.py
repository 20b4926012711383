"""T1 kernel unit tests: every native sm_100a kernel against the plain PyTorch fp32 restatement of the
same op (``ops/reference.py``).  All tests need a B200 and are marked ``gpu``."""
import math

import pytest
import torch

from distributed_training_compare_jax_b200.ops import functional as Fn
from distributed_training_compare_jax_b200.ops import native
from distributed_training_compare_jax_b200.ops import reference as R

pytestmark = pytest.mark.gpu
BF16, F32 = torch.bfloat16, torch.float32


@pytest.fixture(autouse=True)
def _seed():
    torch.manual_seed(1234)
    yield
    torch.cuda.synchronize()


def rel_err(got, want):
    got, want = got.float(), want.float()
    return ((got - want).norm() / want.norm().clamp_min(1e-12)).item()


def randn(*shape, scale=1.0):
    return (torch.randn(*shape, device="cuda") * scale).to(BF16)


def test_library_loads_on_b200():
    native.require_gpu_library()
    assert native.load().dtc_device_arch() >= 100


# ------------------------------------------------------------------------------------------------ GEMM
@pytest.mark.parametrize("a_mn,b_mn", [(False, False), (False, True), (True, True), (True, False)])
@pytest.mark.parametrize("M,N,K,bn", [(128, 128, 64, 128), (256, 384, 512, 128), (512, 256, 320, 64), (384, 512, 256, 256),
                                      (200, 72, 136, 128)])
def test_gemm_majors(a_mn, b_mn, M, N, K, bn):
    if (a_mn and M % 8) or (b_mn and N % 8) or K % 8:
        pytest.skip("leading dimension must be a multiple of 8")
    a = randn(K, M) if a_mn else randn(M, K)
    b = randn(K, N) if b_mn else randn(N, K)
    got = Fn.gemm(a, b, a_mn=a_mn, b_mn=b_mn, block_n=bn)
    want = R.gemm(a, b, a_mn=a_mn, b_mn=b_mn)
    assert rel_err(got, want) < 1e-2


def test_gemm_large_k_pipeline_wrap():
    a, b = randn(256, 4096, scale=0.1), randn(256, 4096, scale=0.1)
    got = Fn.gemm(a, b, stages=3)
    assert rel_err(got, R.gemm(a, b)) < 1e-2


def test_gemm_epilogues():
    M, N, K = 384, 256, 192
    a, w, bias, res = randn(M, K), randn(K, N, scale=0.1), randn(N), randn(M, N)
    want = R.gemm(a, w, b_mn=True, bias=bias)
    assert rel_err(Fn.gemm(a, w, b_mn=True, bias=bias), want) < 1e-2
    assert rel_err(Fn.gemm(a, w, b_mn=True, bias=bias, epi="bias_residual", aux=res),
                   R.gemm(a, w, b_mn=True, bias=bias, epi="bias_residual", aux=res)) < 1e-2
    pre, act = Fn.gemm(a, w, b_mn=True, bias=bias, epi="bias_gelu")
    wpre, wact = R.gemm(a, w, b_mn=True, bias=bias, epi="bias_gelu")
    assert rel_err(pre, wpre) < 1e-2 and rel_err(act, wact) < 1e-2
    f1 = randn(M, N)
    assert rel_err(Fn.gemm(a, w, b_mn=True, epi="dgelu", aux=f1), R.gemm(a, w, b_mn=True, epi="dgelu", aux=f1)) < 1e-2
    got32 = Fn.gemm(a, w, b_mn=True, bias=bias, epi="store_f32", alpha=0.5)
    assert got32.dtype == F32 and rel_err(got32, R.gemm(a, w, b_mn=True, bias=bias, alpha=0.5)) < 1e-2


@pytest.mark.parametrize("splits", [1, 4, 7])
def test_gemm_wgrad_splitk_accumulates(splits):
    Mtok, Kin, Nout = 1024, 256, 384
    x, dy = randn(Mtok, Kin), randn(Mtok, Nout)
    acc = torch.randn(Kin, Nout, device="cuda")
    want = R.gemm(x, dy, a_mn=True, b_mn=True, epi="atomic_f32", out=acc.clone())
    Fn.gemm(x, dy, a_mn=True, b_mn=True, epi="atomic_f32", out=acc, splits=splits)
    assert rel_err(acc, want) < 1e-2


def test_gemm_strided_views_and_vocab_shape():
    # lm_head-like: N not a multiple of the tile, output with a padded leading dimension
    M, K, V, Vp = 256, 128, 1000, 1024
    a, w = randn(M, K), randn(K, Vp, scale=0.2)
    out = torch.zeros(M, Vp, device="cuda", dtype=BF16)
    Fn.gemm(a, w[:, :V], b_mn=True, out=out[:, :V])
    assert rel_err(out[:, :V], R.gemm(a, w[:, :V], b_mn=True)) < 1e-2
    assert out[:, V:].abs().max().item() == 0.0


@pytest.mark.parametrize("impl", [1, 2, 3])
def test_gemm_kernel_variants_agree(impl):
    """one-tile-per-CTA (1), CTA-pair 256x128 (2) and CTA-pair 256x256 (3) kernels against the fp32 reference."""
    with Fn.gemm_impl(impl):
        M, K, N = 1024, 320, 768
        x, w, bias, res = randn(M, K), randn(K, N, scale=0.1), randn(N), randn(M, N)
        assert rel_err(Fn.gemm(x, w, b_mn=True, bias=bias), R.gemm(x, w, b_mn=True, bias=bias)) < 1e-2
        assert rel_err(Fn.gemm(x, w, b_mn=True, bias=bias, epi="bias_residual", aux=res),
                       R.gemm(x, w, b_mn=True, bias=bias, epi="bias_residual", aux=res)) < 1e-2
        pre, act = Fn.gemm(x, w, b_mn=True, bias=bias, epi="bias_gelu")
        wpre, wact = R.gemm(x, w, b_mn=True, bias=bias, epi="bias_gelu")
        assert rel_err(pre, wpre) < 1e-2 and rel_err(act, wact) < 1e-2
        dy, f1 = randn(M, N), randn(M, K)
        assert rel_err(Fn.gemm(dy, w, epi="dgelu", aux=f1), R.gemm(dy, w, epi="dgelu", aux=f1)) < 1e-2   # K-major B
        acc = torch.randn(K, N, device="cuda")
        want = R.gemm(x, dy, a_mn=True, b_mn=True, epi="atomic_f32", out=acc.clone())
        Fn.gemm(x, dy, a_mn=True, b_mn=True, epi="atomic_f32", out=acc, splits=4)
        assert rel_err(acc, want) < 1e-2
        # several tiles per CTA pair (the slab-job pipeline of the pair epilogue crosses tile boundaries)
        M2, K2, N2 = 4096, 192, 2560
        x2, w2, bias2, res2 = randn(M2, K2), randn(K2, N2, scale=0.1), randn(N2), randn(M2, N2)
        assert rel_err(Fn.gemm(x2, w2, b_mn=True, bias=bias2), R.gemm(x2, w2, b_mn=True, bias=bias2)) < 1e-2
        assert rel_err(Fn.gemm(x2, w2, b_mn=True, bias=bias2, epi="bias_residual", aux=res2),
                       R.gemm(x2, w2, b_mn=True, bias=bias2, epi="bias_residual", aux=res2)) < 1e-2
        pre, act = Fn.gemm(x2, w2, b_mn=True, bias=bias2, epi="bias_gelu")
        wpre, wact = R.gemm(x2, w2, b_mn=True, bias=bias2, epi="bias_gelu")
        assert rel_err(pre, wpre) < 1e-2 and rel_err(act, wact) < 1e-2
        dy2, f2 = randn(M2, N2), randn(M2, K2)
        assert rel_err(Fn.gemm(dy2, w2, epi="dgelu", aux=f2), R.gemm(dy2, w2, epi="dgelu", aux=f2)) < 1e-2
        # ragged edges: M, N not multiples of the pair tile
        a2, b2 = randn(600, 200), randn(200, 392, scale=0.1)
        assert rel_err(Fn.gemm(a2, b2, b_mn=True), R.gemm(a2, b2, b_mn=True)) < 1e-2


# ------------------------------------------------------------------------------------------- attention
@pytest.mark.parametrize("B,H,T,hd", [(2, 4, 256, 32), (1, 2, 512, 32), (2, 2, 256, 64), (1, 16, 128, 32)])
def test_flash_attention_fwd_bwd(B, H, T, hd):
    Dl = H * hd
    qkv = randn(B * T, 3 * Dl, scale=1.0)
    out, lse2 = Fn.flash_attn_fwd(qkv, B, H, T, hd)
    qkv32 = qkv.float().requires_grad_(True)
    want = R.attention_from_qkv(qkv32, B, H, T, hd)
    assert rel_err(out, want) < 2e-2
    # log2-domain LSE
    q, k = (qkv32[:, i * Dl:(i + 1) * Dl].reshape(B, T, H, hd).permute(0, 2, 1, 3) for i in range(2))
    s = torch.einsum("bhtd,bhsd->bhts", q, k) * hd ** -0.5
    s = s.masked_fill(~torch.tril(torch.ones(T, T, device="cuda", dtype=torch.bool)), float("-inf"))
    assert torch.allclose(lse2, torch.logsumexp(s, -1) / math.log(2.0), atol=2e-2, rtol=1e-2)
    d_o = randn(B * T, Dl)
    want.backward(d_o.float())
    dqkv = Fn.flash_attn_bwd(qkv, out, d_o, lse2, B, H, T, hd)
    for i, name in enumerate("qkv"):
        e = rel_err(dqkv[:, i * Dl:(i + 1) * Dl], qkv32.grad[:, i * Dl:(i + 1) * Dl])
        assert e < 3e-2, f"d{name} rel err {e}"


# ------------------------------------------------------------------------------------------ layer norm
@pytest.mark.parametrize("M,D", [(64, 512), (1000, 512), (96, 1024), (40, 256)])
def test_layernorm_fwd_bwd(M, D):
    x, g, b = randn(M, D, scale=2.0), (1 + 0.1 * torch.randn(D, device="cuda")).to(BF16), randn(D, scale=0.1)
    y, mean, rstd = Fn.layernorm_fwd(x, g, b)
    wy, wmean, wrstd = R.layernorm_fwd(x, g, b)
    assert rel_err(y, wy) < 1e-2 and rel_err(mean, wmean) < 1e-3 and rel_err(rstd, wrstd) < 1e-3
    dy, dres = randn(M, D), randn(M, D)
    dg = torch.zeros(D, device="cuda")
    db = torch.zeros(D, device="cuda")
    dx = Fn.layernorm_bwd(dy, x, mean, rstd, g, dg, db, dres=dres)
    wdx, wdg, wdb = R.layernorm_bwd(dy, x, wmean, wrstd, g, dres=dres)
    assert rel_err(dx, wdx) < 1e-2 and rel_err(dg, wdg) < 1e-2 and rel_err(db, wdb) < 1e-2


# ------------------------------------------------------------------------------------------- embedding
def test_embed_fwd_bwd_dropout_mask_consistency():
    B, T, D, V, p = 4, 128, 512, 1000, 0.1
    ids = torch.randint(0, V, (B * T,), device="cuda", dtype=torch.int32)
    # strictly positive tables: wte[x] + wpe[t] is never exactly 0, so `h != 0` identifies the kept elements
    wte = (1 + torch.rand(V, D, device="cuda")).to(BF16)
    wpe = (1 + torch.rand(T, D, device="cuda")).to(BF16)
    step = torch.tensor([3], device="cuda", dtype=torch.int32)
    h0 = Fn.embed_fwd(ids, wte, wpe, T, 0.0, 7, step)
    want = wte.float()[ids.long()] + wpe.float().repeat(B, 1)
    assert rel_err(h0, want) < 1e-2
    h = Fn.embed_fwd(ids, wte, wpe, T, p, 7, step)
    keep = h != 0
    frac = 1.0 - keep.float().mean().item()
    assert abs(frac - p) < 0.01
    assert rel_err(h, torch.where(keep, want / (1 - p), torch.zeros_like(want))) < 1e-2
    # a different step gives a different mask; the same step reproduces it
    step2 = torch.tensor([4], device="cuda", dtype=torch.int32)
    assert not torch.equal(Fn.embed_fwd(ids, wte, wpe, T, p, 7, step2) != 0, h != 0)
    assert torch.equal(Fn.embed_fwd(ids, wte, wpe, T, p, 7, step), h)
    # backward regenerates the same mask
    dh = randn(B * T, D)
    dwte = torch.zeros(V, D, device="cuda")
    dwpe = torch.zeros(T, D, device="cuda")
    Fn.embed_bwd(ids, dh, dwte, dwpe, T, p, 7, step)
    g = torch.where(keep, dh.float() / (1 - p), torch.zeros_like(want))
    wdwte = torch.zeros(V, D, device="cuda").index_add_(0, ids.long(), g)
    wdwpe = g.reshape(B, T, D).sum(0)
    assert rel_err(dwte, wdwte) < 1e-3 and rel_err(dwpe, wdwpe) < 1e-3


# --------------------------------------------------------------------------------------- cross entropy
@pytest.mark.parametrize("M,V,ld", [(64, 50258, 50304), (33, 1000, 1000), (16, 4099, 4104)])
def test_cross_entropy_fused(M, V, ld):
    logits = torch.zeros(M, ld, device="cuda", dtype=BF16)
    logits[:, :V] = randn(M, V, scale=3.0)
    labels = torch.randint(0, V, (M,), device="cuda", dtype=torch.int32)
    ref_in = logits[:, :V].clone()
    loss_rows = Fn.ce_fwd_bwd(logits, labels, V, 1.0 / M)
    assert torch.allclose(loss_rows, R.cross_entropy_rows(ref_in, labels), atol=2e-3, rtol=1e-3)
    assert rel_err(logits[:, :V], R.cross_entropy_grad(ref_in, labels, 1.0 / M)) < 1e-2
    assert logits[:, V:].abs().max().item() == 0.0 if ld > V else True


@pytest.mark.parametrize("M,K,V,block_n", [(256, 128, 50258, 256), (512, 64, 1000, 128), (384, 64, 4000, 256)])
def test_lm_head_statistics_epilogue_feeds_the_loss_kernel(M, K, V, block_n):
    """lm_head GEMM with the statistics epilogue + single-pass loss kernel == plain GEMM + two-pass fused CE kernel, and
    both match the fp32 reference (vocabulary not a multiple of anything: pad columns must not enter the statistics)."""
    torch.manual_seed(5)
    Vp = -(-V // 64) * 64
    x, w, bias = randn(M, K), randn(K, Vp, scale=0.3), randn(Vp)
    labels = torch.randint(0, V, (M,), device="cuda", dtype=torch.int32)
    plain = Fn.gemm(x, w, b_mn=True, bias=bias)
    ref_in = plain[:, :V].float()
    n_parts = 2 * (-(-Vp // block_n))
    parts = torch.full((M, n_parts, 2), float("nan"), device="cuda", dtype=F32)
    logits = torch.zeros(M, Vp, device="cuda", dtype=BF16)
    Fn.gemm(x, w, b_mn=True, bias=bias, out=logits, epi="store_stats", block_n=block_n, stats=parts, n_valid=V)
    assert torch.equal(logits, plain)                       # same tiles, same rounding
    # the parts merge to the row's logsumexp over the VALID columns
    m, s = parts[..., 0], parts[..., 1]
    lse = torch.logsumexp(torch.where(s > 0, m + torch.log(s.clamp_min(1e-30)), torch.full_like(m, -float("inf"))), dim=1)
    assert torch.allclose(lse, torch.logsumexp(ref_in, dim=1), atol=2e-3, rtol=1e-4)
    loss_a = Fn.ce_bwd_from_parts(logits, labels, parts, n_parts, V, 1.0 / M)
    loss_b = Fn.ce_fwd_bwd(plain, labels, V, 1.0 / M)
    assert torch.allclose(loss_a, loss_b, atol=1e-3, rtol=1e-4)
    assert torch.allclose(loss_a, R.cross_entropy_rows(ref_in, labels), atol=2e-3, rtol=1e-3)
    assert rel_err(logits[:, :V], R.cross_entropy_grad(ref_in, labels, 1.0 / M)) < 1e-2
    assert logits[:, V:].abs().max().item() == 0.0           # pad columns carry no gradient


def test_cross_entropy_vocab_parallel_matches_fused():
    M, V, n = 48, 1024, 4
    Vl = V // n
    full = randn(M, V, scale=3.0)
    labels = torch.randint(0, V, (M,), device="cuda", dtype=torch.int32)
    shards = [full[:, r * Vl:(r + 1) * Vl].contiguous() for r in range(n)]
    stats = torch.empty(n, M, 3, device="cuda")
    for r in range(n):
        Fn.ce_stats(shards[r], labels, Vl, r * Vl, stats[r])
    lse = torch.empty(M, device="cuda")
    loss_rows = torch.empty(M, device="cuda")
    Fn.ce_combine(stats, n, M, lse, loss_rows)
    assert torch.allclose(loss_rows, R.cross_entropy_rows(full, labels), atol=2e-3, rtol=1e-3)
    for r in range(n):
        Fn.ce_grad(shards[r], labels, lse, Vl, r * Vl, 1.0 / M)
    assert rel_err(torch.cat(shards, 1), R.cross_entropy_grad(full, labels, 1.0 / M)) < 1e-2
    out = torch.zeros(1, device="cuda")
    Fn.reduce_sum(loss_rows, 1.0 / M, out)
    assert abs(out.item() - R.cross_entropy_rows(full, labels).mean().item()) < 1e-3


# ------------------------------------------------------------------------------------------- optimizer
@pytest.mark.parametrize("n", [1 << 20, 12345, 4099])
def test_adamw_clip_matches_optax_semantics(n):
    p = torch.randn(n, device="cuda")
    g = torch.randn(n, device="cuda") * 0.01
    m = torch.zeros(n, device="cuda")
    v = torch.zeros(n, device="cuda")
    pb = torch.empty(n, device="cuda", dtype=BF16)
    step = torch.zeros(1, device="cuda", dtype=torch.int32)
    scalars = torch.zeros(2, device="cuda")
    partials = torch.empty(native.sumsq_num_partials(n), device="cuda")
    rp, rm, rv = p.clone(), m.clone(), v.clone()
    lr, wd, clip = 3e-4, 0.1, 1.0
    for t in range(1, 4):
        gt = g * t
        gcopy = gt.clone()
        Fn.sumsq_partials(gcopy, partials)
        Fn.clip_finalize(partials, partials.numel(), 1.0, clip, scalars, step)
        Fn.adamw(p, m, v, gcopy, pb, scalars, step, lr, 0.9, 0.999, 1e-8, wd)
        norm = gt.norm()
        assert abs(scalars[0].item() - norm.item()) / norm.item() < 1e-4
        R.adamw_step(rp, rm, rv, gt * R.clip_factor(norm, clip), t, lr, wd=wd)
        assert gcopy.abs().max().item() == 0.0  # gradient zeroed for the next step
    assert step.item() == 3
    assert rel_err(p, rp) < 1e-5 and rel_err(m, rm) < 1e-4 and rel_err(v, rv) < 1e-4
    assert rel_err(pb, rp) < 1e-2


def test_colsum_bias_grad():
    M, N = 1000, 1536
    x = randn(M, N)
    out = torch.ones(N, device="cuda")
    Fn.colsum(x, out)
    assert rel_err(out, 1 + x.float().sum(0)) < 1e-3
