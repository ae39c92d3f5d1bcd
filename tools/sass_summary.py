"""Per-kernel counts of the SASS mnemonics that show which hardware path a kernel uses (cuobjdump -sass of the built
library): UTC*MMA = tcgen05.mma, LDTM/STTM = tcgen05.ld/st, UTMALDG/UTMASTG/UTMAREDG = TMA load / store / reduce,
HMMA = mma.sync, LDGSTS = cp.async.  Usage: python tools/sass_summary.py > profiles/rNN_sass_summary.txt"""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
lib = os.path.join(ROOT, "k3m_b200", "libk3m_sm100a.so")
out = subprocess.run(["cuobjdump", "-sass", lib], capture_output=True, text=True).stdout
pats = ["UTCHMMA", "UTCHMMA.2CTA", "LDTM", "STTM", "UTMALDG", "UTMASTG", "UTMAREDG", "UTMAPF", "HMMA", "LDGSTS", "RED", "ATOM", "SYNCS", "UCGABAR"]
cur, counts = None, collections.OrderedDict()
for line in out.splitlines():
    m = re.match(r"\s*Function : (\S+)", line)
    if m:
        cur = m.group(1)
        counts[cur] = collections.Counter()
        continue
    if cur is None:
        continue
    m = re.search(r"^\s+/\*[0-9a-f]+\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_.]+)", line)
    if not m:
        continue
    op = m.group(1)
    for p in pats:
        if op == p or op.startswith(p + "."):
            counts[cur][p] += 1
    if op.startswith("UTCHMMA") and ".2CTA" in op and not op.startswith("UTCHMMA.2CTA"):
        counts[cur]["UTCHMMA.2CTA"] += 1


def demangle(n):
    try:
        return subprocess.run(["cu++filt", n], capture_output=True, text=True).stdout.strip() or n
    except Exception:
        return n


print("# cuobjdump -sass %s: instruction counts per kernel (only kernels with at least one of the listed mnemonics)" % os.path.basename(lib))
print("# " + " ".join(pats))
tot = collections.Counter()
for fn, c in counts.items():
    if not sum(c.values()):
        continue
    name = demangle(fn).replace("(int)", "").replace("(bool)", "").replace("void <unnamed>::", "")
    name = re.sub(r">\(.*", ">", name)
    name = re.sub(r"^([A-Za-z0-9_]+)\(.*", r"\1", name)
    print("%-70s %s" % (name[:70], " ".join("%s=%d" % (p, c[p]) for p in pats if c[p])))
    tot.update(c)
print("TOTAL " + " ".join("%s=%d" % (p, tot[p]) for p in pats if tot[p]))
