"""Symmetric memory for the fused compute+collective kernels.

A :class:`SymmBuffer` is a device tensor allocated with ``torch.distributed._symmetric_memory`` (CUDA VMM
allocations whose handles are exchanged at rendezvous): every rank of the group can address every peer's copy
(``ptrs[r]``) and — on NVSwitch systems with NVLS — one MULTICAST address (``mc_ptr``) bound to all copies, which
is what ``multimem.ld_reduce`` / ``multimem.st`` / ``multimem.red`` in ``csrc/comm.cu`` and the GEMM epilogue use.
:class:`FlagBarrier` is the cross-rank barrier (one int flag per ordered rank pair and slot) launched as a tiny
kernel on the current stream — no host synchronisation, capturable in the step's CUDA graph."""
from __future__ import annotations

import ctypes as C

import torch

from ..ops import native


class SymmBuffer:
    def __init__(self, numel: int, dtype: torch.dtype, device: torch.device, group, zero: bool = True):
        import torch.distributed._symmetric_memory as symm_mem

        self.tensor = symm_mem.empty(numel, dtype=dtype, device=device)
        if zero:
            self.tensor.zero_()
        self.handle = symm_mem.rendezvous(self.tensor, group)
        self.rank, self.world = self.handle.rank, self.handle.world_size
        self.ptrs = [int(p) for p in self.handle.buffer_ptrs]
        self.ptr_array = (C.c_ulonglong * self.world)(*self.ptrs)   # host array handed to the C ABI
        mc = 0
        try:
            if self.handle.has_multicast_support(device.type, device.index if device.index is not None else 0):
                mc = int(self.handle.multicast_ptr)
        except Exception:
            mc = int(getattr(self.handle, "multicast_ptr", 0) or 0)
        self.mc_ptr = mc
        torch.cuda.synchronize()
        self.handle.barrier()


class FlagBarrier:
    N_SLOTS = 16

    def __init__(self, device: torch.device, group):
        import os
        self.buf = SymmBuffer(self.N_SLOTS * 16, torch.int32, device, group)
        self.rank, self.world = self.buf.rank, self.buf.world
        # Protocol stress test (tests/test_multigpu.py): DTC_SYNC_JITTER_NS=<max ns> puts a per-rank, per-call-site,
        # per-step pseudo-random delay in front of every barrier, so ranks arrive in a different order on every replay
        # of the step graph.  `step_counter` (device int that changes between replays) is set by the owner.
        self.jitter_ns = int(os.environ.get("DTC_SYNC_JITTER_NS", "0"))
        self.step_counter = None
        self._calls = 0

    def __call__(self, slot: int) -> None:
        if self.jitter_ns:
            self._calls += 1
            native.call("dtc_jitter", (self.rank * 7919 + self._calls * 104729 + slot) & 0xFFFFFFFF, self.jitter_ns,
                        self.step_counter)
        native.call("dtc_symm_barrier", self.buf.ptr_array, self.rank, self.world, slot)

    @property
    def ptr_array(self):
        return self.buf.ptr_array
