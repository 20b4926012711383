"""Micro-benchmark of the tcgen05 GEMM on the model's shapes: CUDA-event timing, >= 3 warm-up launches, inputs
rotated through enough distinct buffers to exceed the 126 MB L2 (no L2-resident re-reads), TFLOP/s vs the measured
cuBLAS peak in MEASURED_PEAKS.json.  Usage: python tools/gemm_bench.py [--torch] (adds torch.matmul = cuBLAS)."""
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from distributed_training_compare_jax_b200.ops import functional as Fn  # noqa: E402
from distributed_training_compare_jax_b200.ops import native  # noqa: E402

M, D, F, V = 4096, 512, 2048, 50304
for _a in sys.argv:   # --m=512: tokens per GPU of the strong-scaling runs (global batch 8 x 512 tokens over N GPUs)
    if _a.startswith("--m="):
        M = int(_a.split("=")[1])
SHAPES = [  # name, a shape, b shape, a_mn, b_mn, epi, out shape / dtype, splits
    ("qkv fwd", (M, D), (D, 3 * D), False, True, "store", 0),
    ("fc1 fwd+gelu", (M, D), (D, F), False, True, "bias_gelu", 0),
    ("fc2 fwd+res", (M, F), (F, D), False, True, "bias_residual", 0),
    ("out_proj fwd+res", (M, D), (D, D), False, True, "bias_residual", 0),
    ("fc2 dgrad+dgelu", (M, D), (F, D), False, False, "dgelu", 0),
    ("fc1 dgrad", (M, F), (D, F), False, False, "store", 0),
    ("fc1 wgrad", (M, D), (M, F), True, True, "atomic_f32", 8),
    ("fc2 wgrad", (M, F), (M, D), True, True, "atomic_f32", 8),
    ("lm_head fwd", (M, D), (D, V), False, True, "store", 0),
    ("lm_head dgrad", (M, V), (D, V), False, False, "store", 0),
    ("lm_head wgrad", (M, D), (M, V), True, True, "atomic_f32", 1),
]


def run(name, ash, bsh, a_mn, b_mn, epi, splits, impl, iters=20):
    Fn.gemm_impl._tls.current = impl   # (single-threaded tool: set the per-thread variant directly)
    Kd, Md = (ash[0], ash[1]) if a_mn else (ash[1], ash[0])
    Nd = bsh[1] if b_mn else bsh[0]
    bytes_per_set = (ash[0] * ash[1] + bsh[0] * bsh[1] + 2 * Md * Nd) * 2
    nsets = max(2, min(64, (300 << 20) // bytes_per_set + 1))
    sets = []
    for _ in range(nsets):
        a = torch.randn(ash, device="cuda").to(torch.bfloat16)
        b = (torch.randn(bsh, device="cuda") * 0.05).to(torch.bfloat16)
        out = torch.zeros((Md, Nd), device="cuda", dtype=torch.float32 if epi == "atomic_f32" else torch.bfloat16)
        aux = torch.randn((Md, Nd), device="cuda").to(torch.bfloat16) if epi in ("bias_residual", "dgelu") else None
        out2 = torch.empty((Md, Nd), device="cuda", dtype=torch.bfloat16) if epi == "bias_gelu" else None
        bias = torch.randn(Nd, device="cuda").to(torch.bfloat16) if epi.startswith("bias") else None
        sets.append((a, b, out, aux, out2, bias))

    def call(s):
        a, b, out, aux, out2, bias = s
        Fn.gemm(a, b, a_mn=a_mn, b_mn=b_mn, epi=epi, bias=bias, aux=aux, out=out, out2=out2, splits=splits)

    for i in range(3):
        call(sets[i % nsets])
    torch.cuda.synchronize()
    # the host of the GPU box needs ~15 us per launch: time a captured graph so that only device time is measured
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        for i in range(iters):
            call(sets[i % nsets])
    g.replay()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    g.replay()
    e1.record()
    torch.cuda.synchronize()
    us = e0.elapsed_time(e1) * 1e3 / iters
    return us, 2.0 * Md * Nd * Kd / us / 1e6


def main():
    peak = 1408.1
    try:
        peak = json.load(open(os.path.join(os.path.dirname(__file__), "..", "MEASURED_PEAKS.json")))["bf16_tflops_sustained"]
    except Exception:
        pass
    impls = [0, 4, 1, 2, 3]
    for a in sys.argv:
        if a.startswith("--impls="):
            impls = [int(v) for v in a.split("=")[1].split(",")]
    only = [a.split("=")[1] for a in sys.argv if a.startswith("--only=")]
    names = {0: "default", 4: "1-CTA pers", 1: "1-tile/CTA", 2: "pair x128", 3: "pair x256", 5: "pair light"}
    print(f"{'gemm (device time, graph replay)':34s} " + " | ".join(f"{names.get(i, str(i)):>10s} us TFLOP/s frac" for i in impls))
    for name, ash, bsh, a_mn, b_mn, epi, splits in SHAPES:
        if only and not any(o in name for o in only):
            continue
        cells = []
        for i in impls:
            u, t = run(name, ash, bsh, a_mn, b_mn, epi, splits, i)
            cells.append(f"{u:10.1f}    {t:7.1f} {t / peak:4.2f}")
        line = f"{name:34s} " + " | ".join(cells)
        if "--torch" in sys.argv and not a_mn and epi == "store":
            a = torch.randn(ash, device="cuda").to(torch.bfloat16)
            b = torch.randn(bsh, device="cuda").to(torch.bfloat16)
            bb = b if b_mn else b.t()
            o = torch.empty(a.shape[0], bb.shape[1], device="cuda", dtype=torch.bfloat16)
            for _ in range(3):
                torch.matmul(a, bb, out=o)
            torch.cuda.synchronize()
            g = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g):
                for _ in range(20):
                    torch.matmul(a, bb, out=o)
            g.replay()
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            g.replay()
            e1.record()
            torch.cuda.synchronize()
            line += f" | cuBLAS {e0.elapsed_time(e1) * 50:.1f} us (graph, L2-warm)"
        print(line, flush=True)
    Fn.gemm_impl._tls.current = 0


if __name__ == "__main__" and "--grouped" not in sys.argv:
    main()


def bench_grouped():
    """The four weight gradients of one layer as ONE launch: 1-CTA 128x128 tiles vs CTA-pair 256x256 tiles (+ split-K)."""
    K, D, F = 4096, 512, 2048
    n = 6
    sets = []
    for _ in range(n):
        xs = [torch.randn(K, m, device="cuda").to(torch.bfloat16) for m in (F, D, D, D)]
        dys = [(torch.randn(K, c, device="cuda") * 0.02).to(torch.bfloat16) for c in (D, F, D, 3 * D)]
        outs = [torch.zeros(x.shape[1], d.shape[1], device="cuda") for x, d in zip(xs, dys)]
        sets.append((Fn.wgrad_group_args(xs, dys, outs), xs, dys, outs))
    flops = 2.0 * K * (F * D + D * F + D * D + D * 3 * D)
    for pair, splits in ((False, 1), (False, 2), (True, 1), (True, 2), (True, 3), (True, 4)):
        for s in sets[:2]:
            Fn.wgrad_grouped(s[0], splits=splits, pair=pair)
        torch.cuda.synchronize()
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g):
            for s in sets:
                Fn.wgrad_grouped(s[0], splits=splits, pair=pair)
        g.replay()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        g.replay()
        e1.record()
        torch.cuda.synchronize()
        us = e0.elapsed_time(e1) * 1e3 / n
        print(f"grouped wgrad (4 GEMMs of a layer) pair={pair!s:5s} splits={splits}: {us:7.1f} us  {flops / us / 1e6:7.1f} TFLOP/s", flush=True)


if "--grouped" in sys.argv:
    bench_grouped()
