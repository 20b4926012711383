"""Turn an ncu launch list into a per-kernel time budget.
    ncu --metrics gpu__time_duration.sum --clock-control none -c 800 --csv --log-file gpurun_out/launches.csv python bench.py ...
    python tools/ncu_launch_table.py gpurun_out/launches.csv [--steps N] [--md profiles/rX/step_budget.md]
Groups the launches by kernel name (template arguments kept, parameter list dropped), prints count, mean and total device
time, optionally divided by the number of training steps the capture window covered.  Durations measured under ncu are
serialised and inflated for short kernels: use the table for RANKING and shares, not as step time."""
import argparse
import collections
import csv
import re


def load(path: str):
    lines = [l for l in open(path, errors="ignore") if l.startswith('"')]
    rows = []
    for row in csv.DictReader(lines):
        if row.get("Metric Name") != "gpu__time_duration.sum":
            continue
        v = float(row["Metric Value"].replace(",", ""))
        unit = row.get("Metric Unit", "ns")
        v *= {"ns": 1e-3, "nsecond": 1e-3, "us": 1.0, "usecond": 1.0, "ms": 1e3, "msecond": 1e3}.get(unit, 1e-3)
        name = re.sub(r"\(.*", "", row["Kernel Name"]).replace("void ", "").strip()
        rows.append((name, v, row.get("Grid Size", ""), row.get("Block Size", "")))
    return rows


def main() -> None:
    ap = argparse.ArgumentParser()
    ap.add_argument("csv")
    ap.add_argument("--steps", type=float, default=1.0, help="training steps covered by the capture window")
    ap.add_argument("--md", default=None, help="also write the table as markdown")
    args = ap.parse_args()
    rows = load(args.csv)
    agg = collections.OrderedDict()
    for name, us, grid, block in rows:
        a = agg.setdefault(name, [0, 0.0, grid, block])
        a[0] += 1
        a[1] += us
    total = sum(a[1] for a in agg.values())
    out = [f"{len(rows)} launches, {total:.0f} us of kernel time" + (f" = {total / args.steps:.0f} us per step" if args.steps != 1 else ""),
           "", "| kernel | launches/step | mean us | us/step | share |", "|---|---:|---:|---:|---:|"]
    for name, (n, t, grid, block) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        out.append(f"| `{name[:90]}` | {n / args.steps:.1f} | {t / n:.2f} | {t / args.steps:.1f} | {100 * t / total:.1f} % |")
    text = "\n".join(out)
    print(text)
    if args.md:
        with open(args.md, "w") as f:
            f.write(text + "\n")


if __name__ == "__main__":
    main()
