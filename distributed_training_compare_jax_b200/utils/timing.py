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
    """Samples SM clock / throttle reasons of one GPU every ``period`` s while the context is active.

    NVML (``pynvml``, in-process, ≈ 0.1 ms per sample) is used when available so that even a 0.1 s timed region gets
    several samples at the default 10 ms period; otherwise ``nvidia-smi`` is polled (≈ 50 ms per sample)."""

    _REASONS = (("hw_slowdown", 0x8), ("hw_thermal_slowdown", 0x40), ("sw_thermal_slowdown", 0x20), ("sw_power_cap", 0x4))

    def __init__(self, gpu_index: int = 0, period: float = 0.01):
        self.gpu_index, self.period = gpu_index, period
        self.samples: list[list[str]] = []
        self._stop = threading.Event()
        self._t: threading.Thread | None = None
        self.source = "nvidia-smi"
        self._nvml = None
        try:
            import pynvml
            pynvml.nvmlInit()
            # NVML enumerates physical devices: honour CUDA_VISIBLE_DEVICES when it is a plain index list
            vis = os.environ.get("CUDA_VISIBLE_DEVICES", "")
            phys = gpu_index
            if vis and all(v.strip().isdigit() for v in vis.split(",")):
                ids = [int(v) for v in vis.split(",")]
                if gpu_index < len(ids):
                    phys = ids[gpu_index]
            self._handle = pynvml.nvmlDeviceGetHandleByIndex(phys)
            self._max = pynvml.nvmlDeviceGetMaxClockInfo(self._handle, pynvml.NVML_CLOCK_SM)
            self._nvml = pynvml
            self.source = "nvml"
        except Exception:
            self._nvml = None

    def _sample_nvml(self) -> None:
        nv = self._nvml
        sm = nv.nvmlDeviceGetClockInfo(self._handle, nv.NVML_CLOCK_SM)
        try:
            mask = nv.nvmlDeviceGetCurrentClocksEventReasons(self._handle)
        except Exception:
            mask = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self._handle)
        try:
            watts = nv.nvmlDeviceGetPowerUsage(self._handle) / 1e3
        except Exception:
            watts = 0.0
        self.samples.append([str(sm), str(self._max), f"{watts:.1f}"]
                            + ["Active" if mask & bit else "Not Active" for _, bit in self._REASONS])

    def _poll(self) -> None:
        while not self._stop.is_set():
            try:
                if self._nvml is not None:
                    self._sample_nvml()
                else:
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
        sm, mx, watts, reasons = [], [], [], set()
        names = [n for n, _ in self._REASONS]
        for s in self.samples:
            try:
                sm.append(float(s[0]))
                mx.append(float(s[1]))
                watts.append(float(s[2]))
            except (ValueError, IndexError):
                continue
            for name, val in zip(names, s[3:7]):
                if val.lower().startswith("active"):
                    reasons.add(name)
        if not sm:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0, "source": self.source}
        return {"sm_mhz": statistics.median(sm), "sm_max_mhz": max(mx), "reasons": sorted(reasons),
                "samples": len(sm), "power_w_max": max(watts) if watts else None, "source": self.source,
                "period_s": self.period}


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
