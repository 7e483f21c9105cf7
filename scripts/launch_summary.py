"""Summarise an ncu launch list (--metrics gpu__time_duration.sum --csv) of a bench run: the LAST step's launches,
per-kernel total time and share.  usage: python scripts/launch_summary.py launches.csv [last_n_launches]"""
import csv
import sys
from collections import OrderedDict


def main():
    rows = [r for r in csv.reader(open(sys.argv[1])) if len(r) > 14 and r[0].isdigit()]
    lastn = int(sys.argv[2]) if len(sys.argv) > 2 else 0
    if lastn:
        rows = rows[-lastn:]
    agg = OrderedDict()
    for r in rows:
        name = r[4].split("(")[0].replace("void ", "")[:70]
        a = agg.setdefault(name, [0, 0.0])
        a[0] += 1; a[1] += float(r[14]) / 1e3
    tot = sum(a[1] for a in agg.values())
    print("%-70s %5s %10s %6s" % ("kernel", "n", "us", "share"))
    for k, a in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        print("%-70s %5d %10.1f %5.1f%%" % (k, a[0], a[1], 100 * a[1] / tot))
    print("%-70s %5d %10.1f" % ("total", len(rows), tot))


if __name__ == "__main__":
    main()
