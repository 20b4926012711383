"""Build bookkeeping of the native library (ops/build.py): what makes a built libdtc_b200.so stale.  CPU only."""
import importlib
import os

from distributed_training_compare_jax_b200.ops import build


def test_digest_ignores_where_the_tree_is_checked_out():
    # absolute include paths differ between this container and the GPU box (the repo is copied to a scratch path there):
    # they must not be part of the digest, or every library shipped with the snapshot would look stale
    rel = [f for f in build.NVCC_FLAGS if not os.path.isabs(f)]
    assert "-lineinfo" in rel and "arch=compute_100a,code=sm_100a" in rel
    assert all(str(build.REPO_DIR) not in f for f in rel)


def test_extra_defines_are_part_of_the_digest(monkeypatch):
    """DTC_NVCC_DEFINES=DTC_GEMM_TRACE builds the clock-stamp variant of the pair GEMM: a library built with it must not be
    mistaken for the plain one (and vice versa)."""
    plain = build._digest()
    monkeypatch.setenv("DTC_NVCC_DEFINES", "DTC_GEMM_TRACE")
    traced = importlib.reload(build)
    try:
        assert "-DDTC_GEMM_TRACE" in traced.NVCC_FLAGS
        assert traced._digest() != plain
    finally:
        monkeypatch.delenv("DTC_NVCC_DEFINES")
        importlib.reload(build)
    assert build._digest() == plain and "-DDTC_GEMM_TRACE" not in build.NVCC_FLAGS


def test_every_kernel_waits_for_its_predecessor_before_touching_global_memory():
    """Programmatic dependent launch (DTC_PDL=1) lets a kernel start while its predecessor drains; that is only sound if every
    kernel executes griddepcontrol.wait (pdl_wait) — a kernel without it would read its inputs early."""
    import re
    missing = []
    for src in sorted(build.CSRC.glob("*.cu")):
        text = src.read_text()
        for m in re.finditer(r"__global__[^{;]*?(\w+)\s*\(", text):
            start = text.find("{", m.end())
            depth, j = 0, start
            while True:
                if text[j] == "{":
                    depth += 1
                elif text[j] == "}":
                    depth -= 1
                    if depth == 0:
                        break
                j += 1
            body = text[start:j]
            if "pdl_wait" not in body and "jitter_kernel" not in text[m.start():start]:
                missing.append(f"{src.name}:{text[m.start():start].strip()[-60:]}")
    assert not missing, missing
