"""Where the time of one CTA-pair GEMM launch goes (debug build): the first and the last pair's leader CTA stamp clock64 at
the phase boundaries of their first tile.  Build and run with the trace compiled in:

    DTC_NVCC_DEFINES=DTC_GEMM_TRACE python tools/gemm_pair_trace.py [M N K [impl]]

Prints, per traced CTA and averaged over the launches of a graph replay (inputs rotated through > L2), microseconds since
the kernel's first instruction, plus the gap between the end of one launch and the start of the next."""
import ctypes as C
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from distributed_training_compare_jax_b200.ops import functional as Fn  # noqa: E402
from distributed_training_compare_jax_b200.ops import native  # noqa: E402

NAMES = ["entry", "prologue done", "first loads issued", "first stage landed", "last MMA of tile 0 issued",
         "accumulator complete", "  slab 0 in registers", "  slab 0 buffer free", "  slab 0 staged + fenced",
         "last slab store issued", "all stores drained", "all warps done", "TMEM freed"]
IDX = [1, 2, 3, 4, 5, 6, 12, 13, 14, 7, 8, 9, 10]


def main():
    argv = [a for a in sys.argv[1:] if not a.startswith("--")]
    M, N, K = (int(v) for v in argv[:3]) if len(argv) >= 3 else (4096, 1536, 512)
    impl = int(argv[3]) if len(argv) > 3 else 3
    epi = next((a.split("=")[1] for a in sys.argv if a.startswith("--epi=")), "store")
    lib = native.load()
    lib.dtc_gemm_pair_set_trace.argtypes = [C.c_void_p]
    lib.dtc_gemm_pair_set_trace.restype = C.c_int
    buf = torch.zeros(16 + 16 * 4096, dtype=torch.int64, device="cuda")
    iters = 16
    per = (M * K + K * N + 2 * M * N) * 2
    nsets = max(2, min(48, (300 << 20) // per + 1))
    A = [torch.randn(M, K, device="cuda").to(torch.bfloat16) for _ in range(nsets)]
    B = [(torch.randn(K, N, device="cuda") * 0.05).to(torch.bfloat16) for _ in range(nsets)]
    O = [torch.empty(M, N, device="cuda", dtype=torch.bfloat16) for _ in range(nsets)]
    X = [torch.randn(M, N, device="cuda").to(torch.bfloat16) for _ in range(nsets)] if epi == "bias_residual" else None
    bias = torch.randn(N, device="cuda").to(torch.bfloat16) if epi.startswith("bias") else None
    Fn.gemm_impl._tls.current = impl

    def fn(i):
        j = i % nsets
        Fn.gemm(A[j], B[j], a_mn=False, b_mn=True, epi=epi, out=O[j], bias=bias, aux=X[j] if X else None)

    for i in range(3):
        fn(i)
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        for i in range(iters):
            fn(i)
    g.replay()
    torch.cuda.synchronize()
    assert lib.dtc_gemm_pair_set_trace(C.c_void_p(buf.data_ptr())) == 0
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    g.replay()
    e1.record()
    torch.cuda.synchronize()
    lib.dtc_gemm_pair_set_trace(C.c_void_p(0))
    t = buf.cpu()
    n = int(t[0])
    rec = t[16:16 + 16 * n].view(n, 16)
    print(f"M={M} N={N} K={K} impl={impl} epi={epi}: {e0.elapsed_time(e1) * 1e3 / iters:.2f} us per launch, {n} records")
    rec = rec[rec[:, 0].argsort()]
    for which, label in ((0, "first pair"), (None, "last pair")):
        rows = [r for r in rec if (int(r[15]) & 0xffffffff) == (0 if which == 0 else (int(r[15]) >> 32) - 2)]
        if not rows:
            continue
        rows = rows[2:]  # the first launches of the replay start on an idle GPU
        # clock rate from the two timers of the same record
        hz = sum((int(r[10]) - int(r[1])) / max(1, int(r[11]) - int(r[0])) for r in rows) / len(rows)
        print(f"\n{label}: {len(rows)} launches, SM clock ~{hz:.2f} GHz (clock64 / globaltimer)")
        for name, i in zip(NAMES, IDX):
            vals = [(int(r[i]) - int(r[1])) / hz / 1e3 for r in rows if int(r[i]) != 0]
            if vals:
                print(f"  {name:28s} {sum(vals) / len(vals):7.2f} us  (min {min(vals):.2f}, max {max(vals):.2f})")
        gaps = [(int(b[0]) - int(a[11])) / 1e3 for a, b in zip(rows[:-1], rows[1:])]
        if gaps:
            print(f"  end of launch -> entry of the next launch (globaltimer): {sum(gaps) / len(gaps):.2f} us")
        span = [(int(r[11]) - int(r[0])) / 1e3 for r in rows]
        print(f"  entry -> TMEM freed (globaltimer): {sum(span) / len(span):.2f} us")


if __name__ == "__main__":
    main()
