"""Summarise an `ncu --metrics gpu__time_duration.sum --csv` launch list: per-kernel launches, total ns, share."""
import collections
import csv
import sys


def main(path):
    with open(path) as f:
        lines = [l for l in f if not l.startswith("==")]
    agg = collections.OrderedDict()
    for row in csv.DictReader(lines):
        a = agg.setdefault(row["Kernel Name"], [0, 0.0, row["Block Size"], row["Grid Size"]])
        a[0] += 1
        a[1] += float(row["Metric Value"].replace(",", ""))
    tot = sum(a[1] for a in agg.values())
    print("| kernel | launches | block | grid | total us | avg us | share |")
    print("|---|---|---|---|---|---|---|")
    for k, a in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        print("| `%s` | %d | %s | %s | %.1f | %.1f | %.3f |" % (k.split("(")[0], a[0], a[2], a[3], a[1] / 1e3, a[1] / 1e3 / a[0], a[1] / tot))


if __name__ == "__main__":
    main(sys.argv[1])
