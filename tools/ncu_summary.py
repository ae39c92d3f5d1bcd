"""Summarise an `ncu --metrics gpu__time_duration.sum --csv` launch list: per kernel name launches,
total microseconds and share.  Usage: python tools/ncu_summary.py launches.csv [first_launch [n_launches]]"""
import csv
import re
import sys
from collections import OrderedDict


def main():
    path = sys.argv[1]
    first = int(sys.argv[2]) if len(sys.argv) > 2 else 0
    count = int(sys.argv[3]) if len(sys.argv) > 3 else 1 << 30
    rows = []
    with open(path) as f:
        lines = [l for l in f if l.startswith('"')]
    for r in csv.DictReader(lines):
        if r["Metric Name"] != "gpu__time_duration.sum":
            continue
        rows.append((int(r["ID"]), r["Kernel Name"], float(r["Metric Value"].replace(",", "")) / 1e3, r["Grid Size"], r["Block Size"]))
    rows = [r for r in rows if first <= r[0] < first + count]
    agg = OrderedDict()
    for _, name, us, grid, block in rows:
        name = re.sub(r"\(.*\)$", "", name)
        name = re.sub(r"^void ", "", name)
        a = agg.setdefault(name, [0, 0.0])
        a[0] += 1
        a[1] += us
    tot = sum(a[1] for a in agg.values())
    print("# launches %d..%d of %s: %d launches, %.1f us total" % (first, first + len(rows), path, len(rows), tot))
    print("kernel,launches,total_us,share")
    for name, (n, us) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        print('"%s",%d,%.1f,%.3f' % (name, n, us, us / tot))


if __name__ == "__main__":
    main()
