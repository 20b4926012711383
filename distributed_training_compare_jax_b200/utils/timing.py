"""Device-side timing and clock sampling (profiling recipe B200_PROFILING.md).

``DeviceTimer`` brackets a region with CUDA events on the launching stream (host perf_counter on CPU);
``ClockSampler`` polls ``nvidia-smi`` in a background thread during a timed region and summarises SM clocks and
throttle reasons, as required for every reported number; ``l2_flush`` writes a buffer larger than the 126 MB L2."""
from __future__ import annotations

import os
import statistics
import subprocess
import threading
import time

import torch


class DeviceTimer:
    def __init__(self, device: torch.device):
        self.cuda = device.type == "cuda"
        if self.cuda:
            self.e0 = torch.cuda.Event(enable_timing=True)
            self.e1 = torch.cuda.Event(enable_timing=True)
        self.t0 = 0.0

    def start(self) -> None:
        if self.cuda:
            torch.cuda.synchronize()
            self.e0.record()
        self.t0 = time.perf_counter()

    def stop(self) -> float:
        """Returns elapsed milliseconds."""
        if self.cuda:
            self.e1.record()
            torch.cuda.synchronize()
            return float(self.e0.elapsed_time(self.e1))
        return (time.perf_counter() - self.t0) * 1e3


_L2_FLUSH_BUF = None


def l2_flush(device: torch.device, nbytes: int = 256 << 20) -> None:
    """Evict L2 (126 MB on B200) by overwriting a larger scratch buffer."""
    global _L2_FLUSH_BUF
    if device.type != "cuda":
        return
    if _L2_FLUSH_BUF is None or _L2_FLUSH_BUF.numel() < nbytes:
        _L2_FLUSH_BUF = torch.empty(nbytes, dtype=torch.uint8, device=device)
    _L2_FLUSH_BUF.zero_()


_QUERY = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
          "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
          "clocks_event_reasons.sw_power_cap")


class ClockSampler:
    """Samples SM clock / throttle reasons of one GPU every ``period`` s while the context is active."""

    def __init__(self, gpu_index: int = 0, period: float = 0.2):
        self.gpu_index, self.period = gpu_index, period
        self.samples: list[list[str]] = []
        self._stop = threading.Event()
        self._t: threading.Thread | None = None

    def _poll(self) -> None:
        while not self._stop.is_set():
            try:
                out = subprocess.run(["nvidia-smi", f"--id={self.gpu_index}", f"--query-gpu={_QUERY}",
                                      "--format=csv,noheader,nounits"], capture_output=True, text=True, timeout=5)
                if out.returncode == 0 and out.stdout.strip():
                    self.samples.append([c.strip() for c in out.stdout.strip().splitlines()[0].split(",")])
            except Exception:
                pass
            self._stop.wait(self.period)

    def __enter__(self):
        self._t = threading.Thread(target=self._poll, daemon=True)
        self._t.start()
        return self

    def __exit__(self, *exc):
        self._stop.set()
        if self._t is not None:
            self._t.join(timeout=6)

    def summary(self) -> dict:
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for s in self.samples:
            try:
                sm.append(float(s[0]))
                mx.append(float(s[1]))
            except (ValueError, IndexError):
                continue
            for name, val in zip(names, s[3:7]):
                if val.lower().startswith("active"):
                    reasons.add(name)
        if not sm:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        return {"sm_mhz": statistics.median(sm), "sm_max_mhz": max(mx), "reasons": sorted(reasons),
                "samples": len(sm)}


class nvtx_range:
    """``with nvtx_range("forward"):`` — NVTX range around a phase of the step when ``DTC_NVTX=1`` (visible in ncu / nsys
    timelines; the reference has no tracing at all).  A no-op otherwise and on CPU."""
    enabled = os.environ.get("DTC_NVTX") == "1"

    def __init__(self, name: str):
        self.name = name
        self.on = nvtx_range.enabled and torch.cuda.is_available()

    def __enter__(self):
        if self.on:
            torch.cuda.nvtx.range_push(self.name)
        return self

    def __exit__(self, *exc):
        if self.on:
            torch.cuda.nvtx.range_pop()
        return False
