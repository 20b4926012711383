"""ctypes binding of ``libdtc_b200.so`` (declared in ``include/dtc_b200.h``).

The library is plain CUDA C with a C ABI: tensors are passed as raw device pointers and every call is
enqueued on the current torch CUDA stream, so the whole training step can be captured in a CUDA graph.
On a machine with a GPU the library is REQUIRED: a missing / stale ``.so`` raises instead of silently
falling back to PyTorch ops.
"""
from __future__ import annotations

import ctypes as C
import os
from pathlib import Path
from typing import Any

import torch

from . import build as _build

_LIB: C.CDLL | None = None
_LAUNCHES = 0  # number of native kernel-launching calls issued by this process (bench.py reports it)

P, I, L, F, U64, U32 = C.c_void_p, C.c_int, C.c_longlong, C.c_float, C.c_ulonglong, C.c_uint

# name -> argtypes (the trailing cudaStream_t is appended automatically by `call`)
_PROTOS: dict[str, list[Any]] = {
    "dtc_cast_f32_to_bf16": [P, P, L],
    "dtc_cast_f32_to_bf16_2d": [P, L, P, L, I, I],
    "dtc_cast_f32_to_bf16_2d_clear": [P, L, P, L, I, I],
    "dtc_gemm_bf16": [P, L, I, P, L, I, I, I, I, I, F, P, P, L, P, L, P, L, I, I, I],
    "dtc_gemm_bf16_ex": [P, L, I, P, L, I, I, I, I, I, F, P, P, L, P, L, P, L, I, I, I, I],
    "dtc_flash_attn_fwd": [P, P, P, L, P, L, P, I, I, I, I],
    "dtc_flash_attn_bwd": [P, P, P, L, P, L, P, L, P, P, P, L, P, P, L, I, I, I, I],
    "dtc_layernorm_fwd": [P, P, P, P, P, P, I, I, F],
    "dtc_layernorm_bwd": [P, P, P, P, P, P, P, P, P, I, I],
    "dtc_embed_fwd": [P, P, P, P, I, I, I, I, F, U64, P, U32, L],
    "dtc_embed_bwd": [P, P, P, P, I, I, I, I, F, U64, P, U32, L],
    "dtc_ce_fwd_bwd": [P, L, P, P, I, I, F],
    "dtc_ce_stats": [P, L, P, P, I, I, I],
    "dtc_ce_bwd_from_parts": [P, L, P, P, I, P, I, I, F],
    "dtc_ce_combine": [P, I, I, P, P],
    "dtc_ce_grad": [P, L, P, P, I, I, I, F],
    "dtc_reduce_sum": [P, I, F, P, I],
    "dtc_sumsq_partials": [P, L, P],
    "dtc_clip_finalize": [P, I, F, F, P, P],
    "dtc_adamw": [P, P, P, P, P, L, P, P, F, F, F, F, F, F, I],
    "dtc_colsum": [P, L, I, I, P],
    "dtc_symm_barrier": [P, I, I, I],
    "dtc_jitter": [U32, U32, P],
    "dtc_symm_push_f32": [P, U64, P, I, L, L],
    "dtc_sum_slots_f32": [P, L, I, P, L, I],
    "dtc_gemm_bf16_scatter": [P, L, P, L, I, I, I, I, P, I, I, L, I, I],
    "dtc_gemm_wgrad_grouped": [I, P, P, P, P, P, P, I, P, P, F, I, I],
    "dtc_gemm_wgrad_grouped_pair": [I, P, P, P, P, P, P, I, P, P, F, I, I],
    "dtc_colsum_grouped": [I, P, P, P, P, P],
    "dtc_sp_ln_fwd": [P, P, P, I, P, P, P, P, P, U64, P, I, L, I, I, F],
    "dtc_sp_ln_bwd": [P, I, P, P, P, P, P, P, P, U64, P, I, L, I, I],
    "dtc_gemm_bf16_scatter_signal": [P, L, P, L, I, I, I, I, P, I, I, L, I, I, P, P],
    "dtc_sp_ln_fwd_sync": [P, P, P, I, P, P, P, P, P, U64, P, I, L, I, I, F, P, P, I, I, I, I],
    "dtc_sp_ln_bwd_sync": [P, I, P, P, P, P, P, P, P, U64, P, I, L, I, I, P, P, I, I, I, I],
    "dtc_sp_wait": [P, P, I, I],
    "dtc_dp_reduce_scatter": [P, U64, P, I, I, L, L, P],
    "dtc_dp_norm_exchange": [P, I, P, P, I, I, I, F, P, P],
    "dtc_dp_adamw_allgather": [P, P, P, P, U64, P, I, L, L, L, P, P, F, F, F, F, F],
    "dtc_dp_reduce_scatter_ranges": [P, U64, P, I, I, P, P, I, I, P],
    "dtc_dp_adamw_allgather_ranges": [P, P, P, P, U64, P, I, P, P, I, I, P, P, F, F, F, F, F],
}
# launches per call (for the gpu_launches accounting in bench.py)
_LAUNCH_COUNT = {"dtc_flash_attn_bwd": 2}


def lib_path() -> Path:
    return _build.LIB_PATH


def available() -> bool:
    return lib_path().exists()


def load(build_if_missing: bool = True) -> C.CDLL:
    global _LIB
    if _LIB is not None:
        return _LIB
    path = lib_path()
    force = os.environ.get("DTC_REBUILD") == "1"
    if force or not _build.is_fresh():
        # missing OR stale (csrc / header digest differs from the one the .so was built from): never run an old binary
        # against newer sources.  The digest check is cheap (hash of ~10 files); the rebuild needs nvcc.
        have_nvcc = True
        try:
            _build.find_nvcc()
        except RuntimeError:
            have_nvcc = False
        if not build_if_missing or not have_nvcc:
            what = "missing" if not path.exists() else "stale (csrc/ or include/ changed since it was built)"
            raise RuntimeError(f"{path} is {what}: run `python -m distributed_training_compare_jax_b200.ops.build`"
                               + ("" if have_nvcc else " on a machine with nvcc"))
        _build.build(force=force)
    lib = C.CDLL(str(path))
    lib.dtc_last_error.restype = C.c_char_p
    lib.dtc_version.restype = I
    lib.dtc_device_arch.restype = I
    lib.dtc_sumsq_num_partials.restype = I
    lib.dtc_sumsq_num_partials.argtypes = [L]
    lib.dtc_dp_num_partials.restype = I
    lib.dtc_dp_num_partials.argtypes = [L]
    lib.dtc_gemm_set_impl.restype = I
    lib.dtc_gemm_set_impl.argtypes = [I]
    for name, argtypes in _PROTOS.items():
        fn = getattr(lib, name)
        fn.restype = I
        fn.argtypes = [*argtypes, P]
    _LIB = lib
    return lib


def ptr(t: torch.Tensor | None) -> int | None:
    if t is None:
        return None
    return t.data_ptr()


def last_error() -> str:
    return load().dtc_last_error().decode()


def call(name: str, *args: Any) -> None:
    """Invoke an entry point on the current CUDA stream; raises on a non-zero return code."""
    global _LAUNCHES
    lib = load()
    conv = [ptr(a) if isinstance(a, torch.Tensor) else a for a in args]
    stream = torch.cuda.current_stream().cuda_stream
    rc = getattr(lib, name)(*conv, stream)
    if rc != 0:
        raise RuntimeError(f"{name} failed (rc={rc}): {lib.dtc_last_error().decode()}")
    _LAUNCHES += _LAUNCH_COUNT.get(name, 1)


def launches() -> int:
    return _LAUNCHES


def add_launches(n: int) -> None:
    """Account for kernels launched by a CUDA-graph replay (counted once while the graph was captured)."""
    global _LAUNCHES
    _LAUNCHES += n


def sumsq_num_partials(n: int) -> int:
    return int(load().dtc_sumsq_num_partials(n))


def require_gpu_library() -> None:
    """Fail loudly when running on a GPU box without the native library (no silent eager fallback)."""
    if torch.cuda.is_available():
        lib = load(build_if_missing=True)
        arch = lib.dtc_device_arch()
        if arch < 100:
            raise RuntimeError(f"libdtc_b200 targets sm_100a (B200); current device reports sm_{arch}")
