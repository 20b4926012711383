"""Debug: per-k-block clock64 timestamps of CTA 0 of the persistent GEMM (producer: stage free / TMA issued;
MMA thread: stage full / MMAs issued).  python tools/gemm_trace.py [K] [N]"""
import ctypes as C
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from distributed_training_compare_jax_b200.ops import functional as Fn  # noqa: E402
from distributed_training_compare_jax_b200.ops import native  # noqa: E402

K = int(sys.argv[1]) if len(sys.argv) > 1 else 2048
N = int(sys.argv[2]) if len(sys.argv) > 2 else 512
M = 4096
lib = native.load()
lib.dtc_gemm_set_trace.argtypes = [C.c_void_p]
lib.dtc_gemm_set_impl(4)
lib.dtc_gemm_set_dual(int(os.environ.get('DUALBITS', '7')))
a = torch.randn(M, K, device="cuda").to(torch.bfloat16)
b = (torch.randn(K, N, device="cuda") * 0.05).to(torch.bfloat16)
o = torch.empty(M, N, device="cuda", dtype=torch.bfloat16)
for _ in range(3):
    Fn.gemm(a, b, b_mn=True, out=o)
trace = torch.zeros(256 * 4 + 32 * 4, device="cuda", dtype=torch.int64)
lib.dtc_gemm_set_trace(trace.data_ptr())
Fn.gemm(a, b, b_mn=True, out=o)
torch.cuda.synchronize()
lib.dtc_gemm_set_trace(None)
full = trace.cpu()
t = full[:1024].view(256, 4)
te = full[1024:].view(32, 4)
nk = min(K // 64 * ((M // 128) * (N // 128) + 147) // 148, 256)
t0 = int(t[0, 0])
print(f"M={M} N={N} K={K}: k-blocks traced {nk}; columns: stage-free, tma-issued, stage-full, mma-issued (cycles since start)")
prev_full = None
for i in range(min(nk, 40)):
    r = [int(x) - t0 for x in t[i]]
    lat = r[2] - r[1]
    gap = (r[2] - prev_full) if prev_full is not None else 0
    prev_full = r[2]
    print(f"kb {i:3d}: free {r[0]:7d}  issued {r[1]:7d}  full {r[2]:7d}  mma {r[3]:7d} | tma->full {lat:6d}  full-to-full {gap:6d}")

print("epilogue (thread 64) per tile: wait-start, tmem_full seen, compute done, stores done (cycles since start)")
for i in range(8):
    r = [int(x) - t0 for x in te[i]]
    if int(te[i][0]) == 0:
        break
    print(f"tile {i}: start {r[0]:7d}  full {r[1]:7d}  computed {r[2]:7d}  stored {r[3]:7d} | compute {r[2]-r[1]:6d}  store {r[3]-r[2]:6d}")
