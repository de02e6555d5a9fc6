"""Summarise an `ncu --metrics gpu__time_duration.sum --csv` launch list: python tests/launch_summary.py launches.csv [marker]
Prints per-kernel count / total / share over the whole capture and, when `marker` (a kernel-name substring that opens
every step, default arrange_fields) occurs at least twice, the launch-by-launch list of the last complete step."""
import collections
import csv
import sys


def main():
    path = sys.argv[1]
    marker = sys.argv[2] if len(sys.argv) > 2 else "arrange_fields"
    rows = list(csv.reader(open(path)))
    hi = [i for i, r in enumerate(rows) if r and r[0] == "ID"][0]
    head, data = rows[hi], rows[hi + 1:]
    ki, vi, gi = head.index("Kernel Name"), head.index("Metric Value"), head.index("Grid Size")
    L = [(r[ki].split("(")[0].replace("void ", ""), r[gi], float(r[vi].replace(",", "")) / 1e3) for r in data if len(r) > vi]
    agg = collections.OrderedDict()
    for n, _, us in L:
        a = agg.setdefault(n, [0, 0.0])
        a[0] += 1
        a[1] += us
    tot = sum(a[1] for a in agg.values())
    print("%-52s %6s %10s %7s" % ("kernel", "count", "total_us", "share"))
    for n, (c, us) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        print("%-52s %6d %10.1f %6.1f%%" % (n[:52], c, us, 100 * us / tot))
    idx = [i for i, x in enumerate(L) if marker in x[0]]
    if len(idx) >= 2:
        a, b = idx[-2], idx[-1]
        print("\nlast complete step (%d launches, %.1f us serialised):" % (b - a, sum(x[2] for x in L[a:b])))
        for n, g, us in L[a:b]:
            print("  %-50s %-14s %8.2f us" % (n[:50], g, us))


if __name__ == "__main__":
    main()
