"""Fixed-cost probe: time of the GEMM kernels on 4096 x 512 x K for growing K (intercept = launch + prologue +
epilogue, slope = per-k-block cost), L2-warm (same buffers) and L2-cold (rotating buffers)."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from distributed_training_compare_jax_b200.ops import functional as Fn  # noqa: E402
from distributed_training_compare_jax_b200.ops import native  # noqa: E402

M, N = 4096, 512


def t(K, impl, cold):
    global N
    native.load().dtc_gemm_set_impl(impl)
    nset = 24 if cold else 1
    sets = [((torch.randn(M, K, device="cuda")).to(torch.bfloat16), (torch.randn(K, N, device="cuda") * 0.05).to(torch.bfloat16),
             torch.empty(M, N, device="cuda", dtype=torch.bfloat16)) for _ in range(nset)]
    for i in range(3):
        a, b, o = sets[i % nset]
        Fn.gemm(a, b, b_mn=True, out=o)
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()   # the host on the GPU box launches slowly: time a captured graph, not a Python loop
    with torch.cuda.graph(g):
        for i in range(48):
            a, b, o = sets[i % nset]
            Fn.gemm(a, b, b_mn=True, out=o)
    g.replay()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    g.replay()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) * 1e3 / 48


print(f"{'K':>6s} {'persist warm':>13s} {'persist cold':>13s} {'1tile warm':>11s} {'1tile cold':>11s} {'pair128 cold':>13s} {'pair256 cold':>13s}  (us per launch)")
for K in (64, 128, 256, 512, 1024, 2048, 4096):
    print(f"{K:6d} {t(K, 4, False):13.1f} {t(K, 4, True):13.1f} {t(K, 1, False):11.1f} {t(K, 1, True):11.1f} "
          f"{t(K, 2, True):13.1f} {t(K, 3, True):13.1f}", flush=True)
N = 2048
print("N = 2048:")
for K in (64, 512, 2048):
    print(f"{K:6d} {t(K, 4, False):13.1f} {t(K, 4, True):13.1f} {t(K, 1, False):11.1f} {t(K, 1, True):11.1f} "
          f"{t(K, 2, True):13.1f} {t(K, 3, True):13.1f}", flush=True)
# empty-ish kernel launch cost for reference
x = torch.zeros(8, device="cuda")
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(200):
    x.add_(1)
e1.record()
torch.cuda.synchronize()
print(f"tiny torch kernel back-to-back: {e0.elapsed_time(e1) * 5:.1f} us per launch")
native.load().dtc_gemm_set_impl(0)
