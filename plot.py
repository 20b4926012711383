"""Plot tool (component C12; reference: plot.py:6-42).

Reads ``outputs/{dp,tp,pp}/log.csv`` (plus ``outputs/dp_tp`` when present), overlays the loss curves →
``outputs/loss.png`` and draws a per-strategy time bar chart → ``outputs/average_elapsed_time.png``.
Differences: the bar shows the true average step time (last cumulative ``elapsed_time`` /
steps) — the reference sums a cumulative column, which is only proportional to it — and because matplotlib is
not installed on the B200 image the tool always writes ``outputs/summary.csv`` / prints a table and draws the
two PNGs with matplotlib when it imports, otherwise with the dependency-free writer in ``utils/png.py``.
"""
from __future__ import annotations

import csv
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))


def read_log(path: str) -> dict:
    steps, elapsed, loss = [], [], []
    with open(path) as f:
        for row in csv.DictReader(f):
            steps.append(int(row["step"]))
            elapsed.append(float(row["elapsed_time"]))
            loss.append(float(row["loss"]))
    return {"step": steps, "elapsed_time": elapsed, "loss": loss}


def summarize(name: str, log: dict) -> dict:
    n = len(log["step"])
    tail = log["loss"][-min(100, n):]
    return {"strategy": name, "steps": n, "total_time_s": log["elapsed_time"][-1] if n else 0.0,
            "avg_step_time_s": (log["elapsed_time"][-1] / n) if n else 0.0,
            "reference_style_sum_of_cumulative": sum(log["elapsed_time"]),
            "first_loss": log["loss"][0] if n else float("nan"), "final_loss": log["loss"][-1] if n else float("nan"),
            "mean_loss_last_100": sum(tail) / len(tail) if tail else float("nan")}


def main(path_root: str = "outputs/") -> list[dict]:
    logs = {}
    for name in ("dp", "tp", "pp", "dp_tp"):
        p = os.path.join(path_root, name, "log.csv")
        if os.path.exists(p):
            logs[name] = read_log(p)
    if not logs:
        print(f"no log.csv found under {path_root}", file=sys.stderr)
        return []
    rows = [summarize(k, v) for k, v in logs.items()]
    os.makedirs(path_root, exist_ok=True)
    with open(os.path.join(path_root, "summary.csv"), "w", newline="") as f:
        w = csv.DictWriter(f, fieldnames=list(rows[0].keys()))
        w.writeheader()
        w.writerows(rows)
    for r in rows:
        print(f"{r['strategy']:>6}: steps {r['steps']:5d}  avg step {r['avg_step_time_s'] * 1e3:9.3f} ms  "
              f"final loss {r['final_loss']:.4f}  mean(last 100) {r['mean_loss_last_100']:.4f}")
    try:
        import matplotlib
        matplotlib.use("Agg")
        import matplotlib.pyplot as plt
    except Exception:
        # no matplotlib on this machine: draw the same two figures with the built-in PNG writer
        from distributed_training_compare_jax_b200.utils.png import bar_chart, line_chart
        rgb = {"dp": (220, 30, 30), "tp": (30, 160, 30), "pp": (30, 30, 220), "dp_tp": (240, 150, 20)}
        line_chart(os.path.join(path_root, "loss.png"), {k: (v["step"], v["loss"]) for k, v in logs.items()}, rgb, "step",
                   "loss")
        bar_chart(os.path.join(path_root, "average_elapsed_time.png"), {r["strategy"]: r["avg_step_time_s"] for r in rows},
                  rgb, "method", "average step time [sec]")
        print("matplotlib not available: figures drawn with the built-in PNG writer")
        return rows
    colors = {"dp": "red", "tp": "green", "pp": "blue", "dp_tp": "orange"}
    for name, log in logs.items():
        plt.plot(log["step"], log["loss"], label=name, alpha=0.5, color=colors[name])
    plt.xlabel("step")
    plt.ylabel("loss")
    plt.legend()
    plt.savefig(os.path.join(path_root, "loss.png"))
    plt.cla()
    plt.bar([r["strategy"] for r in rows], [r["avg_step_time_s"] for r in rows],
            color=[colors[r["strategy"]] for r in rows])
    plt.xlabel("method")
    plt.ylabel("average step time [sec]")
    plt.savefig(os.path.join(path_root, "average_elapsed_time.png"))
    return rows


if __name__ == "__main__":
    main(*sys.argv[1:2])
