"""Per-kernel opcode census of the built library (cuobjdump -sass): the SASS evidence of the Blackwell-native claim.
    python tools/sass_census.py [scflow_b200/libscflow_b200.so] > profiles/r02_sass_census.txt
tcgen05.mma -> UTC*MMA (.2CTA = cta_group::2), tcgen05.ld -> LDTM, TMA -> UTMALDG / UTMASTG / UBLKCP (bulk copies),
tcgen05.commit -> UTCBAR, mbarrier -> SYNCS, 256-bit global accesses -> LDG/STG.*256; HMMA / HGMMA would be the
legacy mma.sync / wgmma paths (none expected)."""
import collections
import re
import subprocess
import sys

lib = sys.argv[1] if len(sys.argv) > 1 else "scflow_b200/libscflow_b200.so"
out = subprocess.run(["cuobjdump", "-sass", lib], capture_output=True, text=True).stdout
pats = [("UTCHMMA.2CTA", r"\bUTC\w*MMA\.2CTA"), ("UTC*MMA", r"\bUTC\w*MMA\b"), ("LDTM", r"\bLDTM"), ("UTMALDG", r"\bUTMALDG"),
        ("UTMASTG", r"\bUTMASTG"), ("UBLKCP", r"\bUBLKCP"), ("UTCBAR", r"\bUTCBAR"), ("SYNCS", r"\bSYNCS"),
        ("LDG.256", r"\bLDG\S*\.256"), ("STG.256", r"\bSTG\S*\.256"), ("F2FP", r"\bF2FP"), ("HMMA", r"\bHMMA"),
        ("HGMMA", r"\b[HQI]GMMA")]
kern, cur = collections.OrderedDict(), None
for line in out.splitlines():
    m = re.search(r"Function : (\S+)", line)
    if m:
        cur = m.group(1)
        kern[cur] = collections.Counter()
        continue
    if cur is None or "/*" not in line:
        continue
    body = line.split("*/", 1)[1] if "*/" in line else line
    if not re.search(r"^\s+[@A-Z]", body):
        continue
    kern[cur]["instructions"] += 1
    for name, pat in pats:
        if re.search(pat, body):
            kern[cur][name] += 1
def demangle(n):
    try:
        return subprocess.run(["c++filt", n], capture_output=True, text=True).stdout.strip().split("(")[0]
    except Exception:
        return n
cols = ["instructions"] + [p[0] for p in pats]
print(f"SASS opcode census of {lib} (cuobjdump -sass, sm_100a); {len(kern)} kernels")
print(f"{'kernel':72s} " + " ".join(f"{c[:9]:>9s}" for c in cols))
tot = collections.Counter()
for k, c in kern.items():
    tot.update(c)
    print(f"{demangle(k)[-72:]:72s} " + " ".join(f"{c[x]:9d}" for x in cols))
print(f"{'TOTAL':72s} " + " ".join(f"{tot[x]:9d}" for x in cols))
