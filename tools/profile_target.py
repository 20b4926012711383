"""Small single-GPU target for `ncu --set full`: a few launches of the hot kernels on the model's shapes.
Usage: python tools/profile_target.py [gemm|attn|all]"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from distributed_training_compare_jax_b200.ops import functional as Fn  # noqa: E402

what = sys.argv[1] if len(sys.argv) > 1 else "all"
M, D, F, V = 4096, 512, 2048, 50304
bf = torch.bfloat16


def r(*s, scale=1.0):
    return (torch.randn(*s, device="cuda") * scale).to(bf)


impl = int(os.environ.get("DTC_PROFILE_IMPL", "0"))   # GEMM variant (tools/gemm_bench.py: 0 default, 2 / 3 CTA pair)
if impl:
    Fn.gemm_impl._tls.current = impl
if what in ("gemm", "all"):
    x, wqkv, bq = r(M, D), r(D, 3 * D, scale=0.05), r(3 * D)
    w1, b1 = r(D, F, scale=0.05), r(F)
    wlm = r(D, V, scale=0.05)
    res = r(M, D)
    w2, b2, g = r(F, D, scale=0.05), r(D), r(M, F)
    for _ in range(3):
        Fn.gemm(x, wqkv, b_mn=True, bias=bq)                                   # qkv fwd
        Fn.gemm(x, w1, b_mn=True, bias=b1, epi="bias_gelu")                    # fc1 fwd + GELU
        Fn.gemm(g, w2, b_mn=True, bias=b2, epi="bias_residual", aux=res)       # fc2 fwd + residual
        Fn.gemm(x, wlm, b_mn=True)                                             # lm_head fwd
if what in ("attn", "all"):
    B, H, T, hd = 8, 16, 512, 32
    qkv = r(B * T, 3 * H * hd)
    d_o = r(B * T, H * hd)
    for _ in range(3):
        out, lse2 = Fn.flash_attn_fwd(qkv, B, H, T, hd)
        Fn.flash_attn_bwd(qkv, out, d_o, lse2, B, H, T, hd)
torch.cuda.synchronize()
print("profile target done")
if what == "seqpar":
    # One rank's share of the sequence-parallel TP = 8 sublayer boundary, with the 8 "peers" emulated by 8 local buffers
    # (same kernels, same addresses-per-owner logic; the stores stay on this GPU instead of crossing NVLink):
    #   fc2-like scatter GEMM [4096 x 256] x [256 x 512] -> slot `rank` of every owner's partial-sum buffer
    #   reduce (8 slots) + bias + residual + LayerNorm on the 512 owned rows, pushed to the 8 gathered buffers; its backward
    import ctypes as C
    n, rank = 8, 3
    Ms = M // n
    x, w = r(M, F // n), r(F // n, D, scale=0.05)
    parts = torch.zeros(n, n, Ms, D, device="cuda", dtype=bf)            # parts[o] = owner o's [n, Ms, D] slots
    part_ptrs = (C.c_ulonglong * n)(*[parts[o].data_ptr() for o in range(n)])
    gathered = torch.zeros(n, M, D, device="cuda", dtype=bf)              # every "rank"'s gathered operand buffer
    out_ptrs = (C.c_ulonglong * n)(*[gathered[o].data_ptr() for o in range(n)])
    res, bias, gamma, beta = r(Ms, D), r(D), r(D), r(D)
    hsum = torch.empty(Ms, D, device="cuda", dtype=bf)
    mean, rstd = torch.empty(Ms, device="cuda"), torch.empty(Ms, device="cuda")
    dgamma, dbeta = torch.zeros(D, device="cuda"), torch.zeros(D, device="cuda")
    for _ in range(3):
        Fn.gemm_scatter(x, w, part_ptrs, n, rank, D, b_mn=True)
        Fn.sp_ln_fwd(res, bias, parts[rank], n, gamma, beta, hsum, mean, rstd, 0, out_ptrs, n, rank * Ms, 1e-6)
        Fn.sp_ln_bwd(parts[rank], n, hsum, mean, rstd, gamma, res, dgamma, dbeta, 0, out_ptrs, n, rank * Ms)
    torch.cuda.synchronize()
    print("profile target done")
