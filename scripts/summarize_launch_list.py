"""Summarise an ncu launch list (`--metrics gpu__time_duration.sum --csv`) of `bench.py --steps 1 --warmup 3`:
per-kernel totals and shares of the last complete train step (steps are delimited by the q_sample kernel)."""
import collections
import csv
import re
import sys

path = sys.argv[1]
with open(path) as f:
    lines = [l for l in f if not l.startswith("==")]
r = csv.reader(lines)
hdr = next(r)
ci = {h: i for i, h in enumerate(hdr)}
rows = []
for x in r:
    if len(x) < len(hdr) or x[ci["Metric Name"]] != "gpu__time_duration.sum":
        continue
    v = float(x[ci["Metric Value"]].replace(",", ""))
    u = x[ci["Metric Unit"]]
    v = v / 1e3 if u in ("ns", "nsecond") else v * 1e3 if u in ("ms", "msecond") else v
    rows.append((x[ci["Kernel Name"]], v))
idx = [i for i, (k, _) in enumerate(rows) if "q_sample" in k]
lo, hi = idx[-2], idx[-1]
step = rows[lo:hi]
tot = sum(v for _, v in step)
print(f"launch list: {len(rows)} kernels; step slice [{lo}:{hi}] = {len(step)} launches, {tot / 1e3:.2f} ms summed "
      f"(serialised, cold-cache ncu durations)\n")
agg = collections.defaultdict(lambda: [0, 0.0])
for k, v in step:
    k = re.sub(r"<.*", "", k)
    k = re.sub(r"\(.*", "", k)
    ours = k.startswith("b2d::") or "b2d::" in k
    agg[k][0] += 1
    agg[k][1] += v
print("| kernel | launches | ms | share |\n|---|---|---|---|")
for k, (n, v) in sorted(agg.items(), key=lambda kv: -kv[1][1])[:int(sys.argv[2]) if len(sys.argv) > 2 else 40]:
    print(f"| `{k[:90]}` | {n} | {v / 1e3:.3f} | {100 * v / tot:.1f} % |")
ours = sum(v for k, (n, v) in agg.items() if "b2d::" in k)
print(f"\nb2d:: kernels: {ours / 1e3:.2f} ms = {100 * ours / tot:.1f} % of the summed step; "
      f"{sum(n for k, (n, v) in agg.items() if 'b2d::' in k)} launches")
