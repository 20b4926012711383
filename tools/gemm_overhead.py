"""Fixed cost of one GEMM launch: device time (graph replay, inputs rotated through > L2) against K at a fixed M x N, and
against N at a fixed K, for the library's GEMM variants and cuBLAS.  The intercept of the K sweep is what a launch costs
before the first and after the last tensor-core instruction (launch gap, prologue, pipeline fill, epilogue, tail).

    python tools/gemm_overhead.py [--impls=0,3,4]"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from distributed_training_compare_jax_b200.ops import functional as Fn  # noqa: E402


def timed_graph(fn, n):
    for i in range(3):
        fn(i)
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        for i in range(n):
            fn(i)
    g.replay()
    torch.cuda.synchronize()
    best = 1e9
    for _ in range(3):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        g.replay()
        e1.record()
        torch.cuda.synchronize()
        best = min(best, e0.elapsed_time(e1) * 1e3 / n)
    return best


def one(M, N, K, impl, b_mn=True, iters=24):
    per = (M * K + K * N + M * N) * 2
    nsets = max(2, min(48, (300 << 20) // per + 1))
    A = [torch.randn(M, K, device="cuda").to(torch.bfloat16) for _ in range(nsets)]
    Bm = [(torch.randn((K, N) if b_mn else (N, K), device="cuda") * 0.05).to(torch.bfloat16) for _ in range(nsets)]
    O = [torch.empty(M, N, device="cuda", dtype=torch.bfloat16) for _ in range(nsets)]
    if impl == "cublas":
        def fn(i):
            j = i % nsets
            torch.matmul(A[j], Bm[j] if b_mn else Bm[j].t(), out=O[j])
    else:
        Fn.gemm_impl._tls.current = impl

        def fn(i):
            j = i % nsets
            Fn.gemm(A[j], Bm[j], a_mn=False, b_mn=b_mn, epi="store", out=O[j])
    us = timed_graph(fn, iters)
    Fn.gemm_impl._tls.current = 0
    return us


def main():
    impls = [0, 3, 4]
    for a in sys.argv:
        if a.startswith("--impls="):
            impls = [int(v) for v in a.split("=")[1].split(",")]
    cols = impls + ["cublas"]
    names = {0: "default", 4: "1-CTA pers", 1: "1-tile/CTA", 2: "pair x128", 3: "pair x256", "cublas": "cuBLAS"}
    # an empty-ish kernel per graph node: the floor of a dependent kernel node
    x = torch.zeros(32, device="cuda")
    print(f"graph node floor (32-element add_): {timed_graph(lambda i: x.add_(1.0), 64):.2f} us")
    M, N = 4096, 1536
    print(f"\nK sweep at M={M} N={N} (us per launch)\n{'K':>6s} " + " ".join(f"{names[c]:>11s}" for c in cols))
    for K in (64, 128, 256, 512, 1024, 2048, 4096):
        print(f"{K:6d} " + " ".join(f"{one(M, N, K, c):11.2f}" for c in cols), flush=True)
    K = 512
    print(f"\nN sweep at M={M} K={K} (us per launch)\n{'N':>6s} " + " ".join(f"{names[c]:>11s}" for c in cols))
    for N in (256, 512, 1024, 1536, 2048, 3072, 4096):
        print(f"{N:6d} " + " ".join(f"{one(M, N, K, c):11.2f}" for c in cols), flush=True)
    print(f"\nK-major B (dgrad orientation), M={M} N=512 (us per launch)\n{'K':>6s} " + " ".join(f"{names[c]:>11s}" for c in cols))
    for K in (512, 2048):
        print(f"{K:6d} " + " ".join(f"{one(M, 512, K, c, b_mn=False):11.2f}" for c in cols), flush=True)


if __name__ == "__main__":
    main()
