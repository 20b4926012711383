"""Device time of the LayerNorm kernels at the model's shape (graph replay, operands rotated through > L2).
    python tools/ln_bench.py [M D]          (DTC_LN_BWD_V1=1: the 8-warp backward kernel)"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from distributed_training_compare_jax_b200.ops import functional as Fn  # noqa: E402

M, D = (int(sys.argv[1]), int(sys.argv[2])) if len(sys.argv) > 2 else (4096, 512)
bf = torch.bfloat16
n = 40  # 40 x 4 tensors x 4 MB > L2
xs = [torch.randn(M, D, device="cuda").to(bf) for _ in range(n)]
dys = [torch.randn(M, D, device="cuda").to(bf) for _ in range(n)]
drs = [torch.randn(M, D, device="cuda").to(bf) for _ in range(n)]
outs = [torch.empty(M, D, device="cuda", dtype=bf) for _ in range(n)]
gamma, beta = torch.randn(D, device="cuda").to(bf), torch.randn(D, device="cuda").to(bf)
mean, rstd = torch.empty(M, device="cuda"), torch.empty(M, device="cuda")
dg, db = torch.zeros(D, device="cuda"), torch.zeros(D, device="cuda")


def timed(fn):
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


mean.normal_()
rstd.uniform_(0.5, 1.5)
t_bwd = timed(lambda i: Fn.layernorm_bwd(dys[i % n], xs[i % n], mean, rstd, gamma, dg, db, dres=drs[i % n], out=outs[i % n]))
t_bwd_nores = timed(lambda i: Fn.layernorm_bwd(dys[i % n], xs[i % n], mean, rstd, gamma, dg, db, out=outs[i % n]))
t_fwd = timed(lambda i: Fn.layernorm_fwd(xs[i % n], gamma, beta, out=outs[i % n], mean=mean, rstd=rstd))
print(f"layernorm_fwd M={M} D={D}: {t_fwd:.2f} us ({2 * M * D * 2 / t_fwd / 1e6:.2f} TB/s)")
byt = 4 * M * D * 2
print(f"layernorm_bwd M={M} D={D} {'(8-warp kernel)' if os.environ.get('DTC_LN_BWD_V1') else ''}: {t_bwd:.2f} us with residual "
      f"({byt / t_bwd / 1e6:.2f} TB/s), {t_bwd_nores:.2f} us without")
