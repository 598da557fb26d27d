"""Summarise an `ncu --metrics gpu__time_duration.sum --csv` launch list per kernel name."""
import collections, csv, re, sys
lines = [l for l in open(sys.argv[1]) if not l.startswith("==")]
agg, tot = collections.defaultdict(lambda: [0, 0.0]), 0.0
for row in csv.DictReader(lines):
    if row.get("Metric Name") != "gpu__time_duration.sum":
        continue
    v = float(row["Metric Value"].replace(",", ""))
    v = v / 1e3 if row["Metric Unit"] == "ns" else v * 1e3 if row["Metric Unit"] == "ms" else v
    name = re.sub(r"<.*", "", row["Kernel Name"])[:78]
    agg[name][0] += 1; agg[name][1] += v; tot += v
print(f"total {tot:.0f} us over {sum(a[0] for a in agg.values())} launches")
top = int(sys.argv[2]) if len(sys.argv) > 2 else 30
for k, (c, t) in sorted(agg.items(), key=lambda kv: -kv[1][1])[:top]:
    print(f"{t:10.0f} us {100*t/tot:5.1f}%  x{c:4d}  {k}")
