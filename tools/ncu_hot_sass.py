"""Top SASS instructions by warp-stall samples from `ncu -i rep --page source --csv` (needs -lineinfo builds).
A report with several launches holds one section per launch; pick one with --section K (default: the last).
    ncu -i gpurun_out/prof.ncu-rep --page source --csv > /tmp/src.csv
    python tools/ncu_hot_sass.py /tmp/src.csv [N] [--section K]"""
import csv
import io
import sys

text = open(sys.argv[1]).read().split("\n")
starts = [i for i, l in enumerate(text) if l.startswith('"Kernel Name"')] + [len(text)]
sec = int(sys.argv[sys.argv.index("--section") + 1]) if "--section" in sys.argv else len(starts) - 2
chunk = "\n".join(text[starts[sec]:starts[sec + 1]])
rows = list(csv.reader(io.StringIO(chunk)))
n = int(sys.argv[2]) if len(sys.argv) > 2 and sys.argv[2].isdigit() else 25
hdr_i = next(i for i, r in enumerate(rows) if "Source" in r and "# Samples" in r)
hdr = rows[hdr_i]
col = {h: i for i, h in enumerate(hdr)}
data = [r for r in rows[hdr_i + 1:] if len(r) == len(hdr)]
stalls = [h for h in hdr if h.startswith("stall_") and "Not Issued" not in h]
total = sum(int(r[col["# Samples"]] or 0) for r in data)
print(f"section {sec} of {len(starts) - 1}: {rows[0][1][:90]}")
print(f"{len(data)} instructions, {total} samples")
agg = {s: sum(int(r[col[s]] or 0) for r in data) for s in stalls}
print("overall: " + ", ".join(f"{k[6:]} {v * 100 // max(total, 1)}%" for k, v in sorted(agg.items(), key=lambda kv: -kv[1])[:8]))
idx = {id(r): i for i, r in enumerate(data)}
for r in sorted(data, key=lambda r: -int(r[col["# Samples"]] or 0))[:n]:
    s = int(r[col["# Samples"]] or 0)
    top = sorted(((int(r[col[k]] or 0), k[6:]) for k in stalls), reverse=True)[:2]
    print(f"{s:6d} {s * 100 / max(total, 1):5.1f}%  #{idx[id(r)]:5d}  exec {r[col['Instructions Executed']]:>8s}  "
          f"{r[col['Source']].strip()[:70]:70s}  " + ", ".join(f"{k} {v}" for v, k in top if v))
