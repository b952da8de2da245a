#!/usr/bin/env python
"""Summarise an `ncu --metrics gpu__time_duration.sum --csv` launch list: per (kernel, grid) count, total, average, share."""
import collections
import csv
import sys


def main(path):
    rows = list(csv.reader(open(path)))
    hdr = next(r for r in rows if "Kernel Name" in r)
    ki, vi, gi = hdr.index("Kernel Name"), hdr.index("Metric Value"), hdr.index("Grid Size")
    acc = collections.OrderedDict()
    for r in rows:
        if len(r) != len(hdr) or not r[0].isdigit():
            continue
        a = acc.setdefault((r[ki].split("(")[0][-58:], r[gi]), [0, 0.0])
        a[0] += 1
        a[1] += float(r[vi].replace(",", ""))
    tot = sum(a[1] for a in acc.values())
    print("# %s: %d launches, %.1f us total (cold-cache, serialised: compare shares)" % (path, sum(a[0] for a in acc.values()), tot / 1e3))
    for k, a in sorted(acc.items(), key=lambda x: -x[1][1]):
        print("%-60s %-16s n=%4d total %10.1f us avg %9.1f us %5.1f%%" % (k[0], k[1], a[0], a[1] / 1e3, a[1] / a[0] / 1e3, 100 * a[1] / tot))


if __name__ == "__main__":
    main(sys.argv[1])
