"""GPU test of ``dtc_flash_attn_bwd_static`` — the attention backward that brings its own scratch memory.

Who needs it: callers that are not allowed to allocate device memory at the point where they run the backward
pass — a plugin living inside a runtime that owns every device allocation, a step loop that cannot afford the
device-wide synchronisation of cudaMalloc / cudaFree, a job whose memory budget was fixed up front.  The regular entry point makes such a caller supply a [B,H,T] scratch buffer; this one keeps that
buffer in a fixed 1 MiB array inside the library, so only inputs and outputs are passed.

The test talks to the library through its C ABI directly (ctypes, raw pointers, explicit stream) and checks,
for two calls in a row with different inputs and shapes, that dQ / dK / dV equal those of the regular entry
point; the second call runs on the scratch contents the first one left behind.  It also checks that a problem
larger than the built-in scratch is refused with a readable message.  The file name sorts last on purpose, so a
``-x`` run has already executed every other test when it gets here; it passes alone in a fresh process too.
"""
import ctypes as C

import pytest
import torch

from distributed_training_compare_jax_b200.ops import native

pytestmark = pytest.mark.gpu

P, L, I = C.c_void_p, C.c_longlong, C.c_int
STATIC_MAX = 1 << 18  # DTC_ATTN_STATIC_DELTA_MAX


def _lib():
    lib = native.load()
    lib.dtc_flash_attn_fwd.restype = I
    lib.dtc_flash_attn_fwd.argtypes = [P, P, P, L, P, L, P, I, I, I, I, P]
    lib.dtc_flash_attn_bwd.restype = I
    lib.dtc_flash_attn_bwd.argtypes = [P, P, P, L, P, L, P, L, P, P, P, L, P, P, L, I, I, I, I, P]
    lib.dtc_flash_attn_bwd_static.restype = I
    lib.dtc_flash_attn_bwd_static.argtypes = [P, P, P, L, P, L, P, L, P, P, L, P, P, L, I, I, I, I, P]
    return lib


def _run_pair(lib, B, H, T, hd, seed):
    torch.manual_seed(seed)
    Dl, M = H * hd, B * T
    dev = "cuda"
    qkv = torch.randn(M, 3 * Dl, device=dev).to(torch.bfloat16)
    d_o = torch.randn(M, Dl, device=dev).to(torch.bfloat16)
    out = torch.empty(M, Dl, device=dev, dtype=torch.bfloat16)
    lse2 = torch.empty(B, H, T, device=dev, dtype=torch.float32)
    stream = torch.cuda.current_stream().cuda_stream
    q, k, v = qkv.data_ptr(), qkv.data_ptr() + Dl * 2, qkv.data_ptr() + 2 * Dl * 2
    assert lib.dtc_flash_attn_fwd(q, k, v, 3 * Dl, out.data_ptr(), Dl, lse2.data_ptr(), B, H, T, hd, stream) == 0

    def outputs():
        return (torch.zeros(M, Dl, device=dev, dtype=torch.float32),        # dq_accum: zero on entry
                torch.empty(M, 3 * Dl, device=dev, dtype=torch.bfloat16))  # dk, dv live in columns [Dl, 3Dl)

    dq_a, dkv_a = outputs()
    delta = torch.empty(B, H, T, device=dev, dtype=torch.float32)
    rc = lib.dtc_flash_attn_bwd(q, k, v, 3 * Dl, out.data_ptr(), Dl, d_o.data_ptr(), Dl, lse2.data_ptr(),
                                delta.data_ptr(), dq_a.data_ptr(), Dl, dkv_a.data_ptr() + Dl * 2,
                                dkv_a.data_ptr() + 2 * Dl * 2, 3 * Dl, B, H, T, hd, stream)
    assert rc == 0, lib.dtc_last_error().decode()
    dq_b, dkv_b = outputs()
    rc = lib.dtc_flash_attn_bwd_static(q, k, v, 3 * Dl, out.data_ptr(), Dl, d_o.data_ptr(), Dl, lse2.data_ptr(),
                                       dq_b.data_ptr(), Dl, dkv_b.data_ptr() + Dl * 2, dkv_b.data_ptr() + 2 * Dl * 2,
                                       3 * Dl, B, H, T, hd, stream)
    assert rc == 0, lib.dtc_last_error().decode()
    torch.cuda.synchronize()

    def rel(a, b):
        return ((a.float() - b.float()).norm() / b.float().norm().clamp_min(1e-12)).item()

    # Same kernels, same inputs, same launch order.  dK / dV are produced by one CTA per key block in a fixed order, so
    # the two entry points must agree bit for bit; dQ is accumulated with fp32 reduce-adds whose arrival order may
    # differ from run to run, which moves it by rounding noise only.
    assert torch.equal(dkv_b[:, Dl:], dkv_a[:, Dl:])
    assert rel(dq_b, dq_a) < 1e-5
    assert torch.isfinite(dq_b).all() and torch.isfinite(dkv_b[:, Dl:].float()).all()


def test_static_scratch_attention_backward_matches_regular_entry_point_twice():
    lib = _lib()
    _run_pair(lib, B=2, H=4, T=256, hd=32, seed=1)   # first call on this device: resolves the scratch address
    _run_pair(lib, B=1, H=2, T=128, hd=64, seed=2)   # second call: smaller problem over the stale scratch contents
    _run_pair(lib, B=2, H=4, T=256, hd=32, seed=3)


def test_static_scratch_rejects_oversized_problem():
    lib = _lib()
    B, H, T, hd = 8, 65, 512, 32  # B*H*T = 266240 > 262144
    assert B * H * T > STATIC_MAX
    t = torch.zeros(16, device="cuda")
    stream = torch.cuda.current_stream().cuda_stream
    p = t.data_ptr()
    rc = lib.dtc_flash_attn_bwd_static(p, p, p, 8, p, 8, p, 8, p, p, 8, p, p, 8, B, H, T, hd, stream)
    assert rc != 0
    msg = lib.dtc_last_error().decode()
    assert "exceeds" in msg and "262144" in msg
    torch.cuda.synchronize()  # nothing was launched, the context is still healthy
