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
    native.load().dtc_gemm_set_impl(impl)
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
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(iters):
        call(sets[i % nsets])
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
    print(f"{'gemm':22s} {'1-CTA persistent us':>20s} {'TFLOP/s':>8s} {'frac':>5s} | {'pair 256x128 us':>15s} {'TFLOP/s':>8s} | "
          f"{'pair 256x256 us':>15s} {'TFLOP/s':>8s} {'frac':>5s}")
    for name, ash, bsh, a_mn, b_mn, epi, splits in SHAPES:
        u2, t2 = run(name, ash, bsh, a_mn, b_mn, epi, splits, 0)
        u1, t1 = run(name, ash, bsh, a_mn, b_mn, epi, splits, 2)
        u3, t3 = run(name, ash, bsh, a_mn, b_mn, epi, splits, 3)
        line = (f"{name:22s} {u2:20.1f} {t2:8.1f} {t2 / peak:5.2f} | {u1:15.1f} {t1:8.1f} | {u3:15.1f} {t3:8.1f} "
                f"{t3 / peak:5.2f}")
        if "--torch" in sys.argv and not a_mn and epi == "store":
            a = torch.randn(ash, device="cuda").to(torch.bfloat16)
            b = torch.randn(bsh, device="cuda").to(torch.bfloat16)
            bb = b if b_mn else b.t()
            for _ in range(3):
                torch.matmul(a, bb)
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(20):
                torch.matmul(a, bb)
            e1.record()
            torch.cuda.synchronize()
            line += f" | cuBLAS {e0.elapsed_time(e1) * 50:.1f} us (L2-warm)"
        print(line, flush=True)
    native.load().dtc_gemm_set_impl(0)


if __name__ == "__main__":
    main()
