"""Summarise an ncu launch list (gpu__time_duration.sum CSV) per kernel: count, total ms, share."""
import collections, csv, re, sys
src, dst = sys.argv[1], sys.argv[2]
lines = [l for l in open(src) if not l.startswith("==")]
agg = collections.defaultdict(lambda: [0, 0.0])
for row in csv.DictReader(lines):
    if row.get("Metric Name") != "gpu__time_duration.sum":
        continue
    k = re.sub(r"\(.*", "", row["Kernel Name"])[:90]
    v = float(row["Metric Value"].replace(",", ""))
    v *= {"ns": 1e-6, "us": 1e-3, "ms": 1.0, "s": 1e3}.get(row["Metric Unit"], 1e-6)
    agg[k][0] += 1
    agg[k][1] += v
tot = sum(v[1] for v in agg.values())
with open(dst, "w") as f:
    f.write(f"# source: {src}; {sum(v[0] for v in agg.values())} launches, {tot:.2f} ms of kernel time (cold-cache, serialised)\n")
    f.write("kernel,launches,total_ms,share_pct\n")
    for k, v in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        f.write(f"\"{k}\",{v[0]},{v[1]:.3f},{100 * v[1] / tot:.2f}\n")
