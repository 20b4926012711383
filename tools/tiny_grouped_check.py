"""Seconds-long check of the grouped CTA-pair weight-gradient kernel on tensor-parallel shard shapes, without importing torch
(driver API through cuda-python + ctypes): out_p = X_p^T dY_p for all-ones operands must be K everywhere, and rows / columns
past the problem's extent must stay untouched.  Problems: [64 x 512] (fewer rows than a TMA box), [512 x 192] (ragged N)."""
import ctypes as C
import os
import sys
import time

import numpy as np
from cuda.bindings import driver as cu

t0 = time.time()
here = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
lib = C.CDLL(os.path.join(here, "distributed_training_compare_jax_b200", "lib", "libdtc_b200.so"))
lib.dtc_last_error.restype = C.c_char_p


def ck(r):
    if isinstance(r, tuple):
        err, *rest = r
        assert int(err) == 0, err
        return rest[0] if rest else None
    assert int(r) == 0, r


ck(cu.cuInit(0))
dev = ck(cu.cuDeviceGet(0))
ctx = ck(cu.cuDevicePrimaryCtxRetain(dev))
ck(cu.cuCtxSetCurrent(ctx))
K = 256
probs = [(64, 512, 128), (512, 192, 512)]   # (M_p, N_p, rows allocated for the output)
ONE = np.uint16(0x3F80)
A, B, O, outs = [], [], [], []
for M, N, rows in probs:
    x = np.full((K, M), ONE, np.uint16)
    dy = np.full((K, N), ONE, np.uint16)
    dx = ck(cu.cuMemAlloc(x.nbytes))
    ddy = ck(cu.cuMemAlloc(dy.nbytes))
    out = np.zeros((rows, 512), np.float32)
    do = ck(cu.cuMemAlloc(out.nbytes))
    ck(cu.cuMemcpyHtoD(dx, x, x.nbytes))
    ck(cu.cuMemcpyHtoD(ddy, dy, dy.nbytes))
    ck(cu.cuMemcpyHtoD(do, out, out.nbytes))
    A.append(int(dx)); B.append(int(ddy)); O.append(int(do)); outs.append(out)
n = len(probs)
u64, i64, i32 = C.c_ulonglong * n, C.c_longlong * n, C.c_int * n
rc = lib.dtc_gemm_wgrad_grouped_pair(C.c_int(n), u64(*A), i64(*[p[0] for p in probs]), u64(*B), i64(*[p[1] for p in probs]),
                                     i32(*[p[0] for p in probs]), i32(*[p[1] for p in probs]), C.c_int(K), u64(*O),
                                     i64(*[512] * n), C.c_float(1.0), C.c_int(1), C.c_int(0), C.c_void_p(0))
if rc != 0:
    print("FAILED rc", rc, lib.dtc_last_error().decode())
    sys.exit(1)
ck(cu.cuCtxSynchronize())
ok = True
for (M, N, rows), out, do in zip(probs, outs, O):
    ck(cu.cuMemcpyDtoH(out, do, out.nbytes))
    good = bool((out[:M, :N] == K).all()) and bool((out[M:, :] == 0).all()) and bool((out[:, N:] == 0).all())
    print(f"[{M} x {N}] inside == {K}: {bool((out[:M, :N] == K).all())}, outside untouched: "
          f"{bool((out[M:, :] == 0).all() and (out[:, N:] == 0).all())}")
    ok &= good
print("OK" if ok else "MISMATCH", f"{time.time() - t0:.1f} s")
sys.exit(0 if ok else 1)
