"""Metrics / logging (reference: main.py:36, train/train.py:63,72,87-102).

Keeps the reference's exact stdout format and the ``log.csv`` schema (columns ``step, elapsed_time, loss`` —
step 0-based, elapsed_time cumulative host seconds since the start of the timed loop) so tooling written for
the reference (``plot.py``) works unchanged, and adds a ``metrics.json`` with device-timed step time (max over
ranks), tokens/s and run metadata that the reference does not record."""
from __future__ import annotations

import csv
import json
import os
import time

import numpy as np


class RunLogger:
    def __init__(self, output_dir: str, log_every: int, is_main: bool = True):
        self.output_dir, self.log_every, self.is_main = output_dir, log_every, is_main
        self.running: list[float] = []
        self.loss_history: list[float] = []
        self.elapsed: list[float] = []
        self.start = 0.0

    def banner(self, strategy: str, n: int) -> None:
        if self.is_main:
            print(f"Running `{strategy}` on {n} devices.", flush=True)

    def say(self, msg: str) -> None:
        if self.is_main:
            print(msg, flush=True)

    def start_timing(self) -> None:
        self.start = time.perf_counter()

    def record(self, step: int, loss: float) -> None:
        now = time.perf_counter()
        self.running.append(loss)
        self.loss_history.append(loss)
        self.elapsed.append(now - self.start)
        if step % self.log_every == 0:
            self.say(f"Step: {step} | Avg loss: {np.mean(self.running):.4f} | "
                     f"Average step time: {(now - self.start) / step:.4f}")
            self.running = []

    def finish(self, extra: dict | None = None) -> None:
        end = time.perf_counter()
        self.say(f"Total time: {end - self.start}")
        self.say("End")
        if not self.is_main:
            return
        os.makedirs(self.output_dir, exist_ok=True)
        with open(os.path.join(self.output_dir, "log.csv"), "w", newline="") as f:
            w = csv.writer(f)
            w.writerow(["step", "elapsed_time", "loss"])
            for i, (t, l) in enumerate(zip(self.elapsed, self.loss_history)):
                w.writerow([i, t, l])
        if extra is not None:
            with open(os.path.join(self.output_dir, "metrics.json"), "w") as f:
                json.dump(extra, f, indent=1)
