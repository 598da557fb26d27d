"""Condense an `ncu --page source --csv` export (SASS view) into regions of equal execution count.
usage: python tools/ncu_source.py gpurun_out/x_source.csv [--dump lo hi]"""
import csv, sys
rows = list(csv.reader(open(sys.argv[1])))
hdr = next(r for r in rows if r and r[0] == "Address")
ix = {h: i for i, h in enumerate(hdr)}
data = [r for r in rows if len(r) == len(hdr) and r[0] != "Address"]
first, seen = [], set()
for r in data:
    if r[0] in seen:
        break
    seen.add(r[0]); first.append(r)
E = lambda r: int(r[ix["Instructions Executed"]] or 0)
S = lambda r: int(r[ix["# Samples"]] or 0)
tot = sum(E(r) for r in first); stot = sum(S(r) for r in first)
print("warp instructions executed:", tot, " SASS lines:", len(first), " samples:", stot)
if "--dump" in sys.argv:
    lo, hi = int(sys.argv[sys.argv.index("--dump") + 1]), int(sys.argv[sys.argv.index("--dump") + 2])
    for i in range(lo, hi + 1):
        r = first[i]
        print(f"{i:4d} {E(r):9d} s={S(r):5d} shW={r[ix['L1 Wavefronts Shared']]:>8s} {r[ix['Source']][:90]}")
    sys.exit()
cur, start, acc, sm, out = None, 0, 0, 0, []
for i, r in enumerate(first):
    e = E(r)
    if cur is None or abs(e - cur) > 0.02 * max(cur, 1):
        if cur is not None:
            out.append((start, i - 1, cur, acc, sm))
        cur, start, acc, sm = e, i, 0, 0
    acc += e; sm += S(r)
out.append((start, len(first) - 1, cur, acc, sm))
for o in out:
    if o[3] > 0.004 * tot or o[4] > 0.01 * stot:
        print(f"sass[{o[0]:4d}-{o[1]:4d}] n={o[1]-o[0]+1:4d} exec/instr={o[2]:9d} total={o[3]/tot*100:5.1f}% samples={o[4]/max(stot,1)*100:5.1f}%")
