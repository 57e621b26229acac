"""Summarise an `ncu --metrics gpu__time_duration.sum --csv` launch list: per (kernel, grid) totals of one step."""
import collections
import csv
import sys


def main(path, first, count):
    with open(path) as f:
        lines = [l for l in f if l.startswith('"')]
    rows = list(csv.DictReader(lines))
    step = rows[first:first + count] if count else rows[first:]
    agg = collections.OrderedDict()
    tot = 0.0
    for x in step:
        d = float(x["Metric Value"].replace(",", "")) / 1e3
        tot += d
        k = (x["Kernel Name"].split("(")[0][-28:], x["Grid Size"], x["Block Size"])
        a = agg.setdefault(k, [0, 0.0])
        a[0] += 1
        a[1] += d
    print(f"launches {len(step)} of {len(rows)}, total {tot:.1f} us")
    for k, v in sorted(agg.items(), key=lambda kv: -kv[1][1])[:30]:
        print(f"{v[1]:10.1f} us {100 * v[1] / tot:5.1f}%  x{v[0]:<3d} {k[0]:30s} grid {k[1]} block {k[2]}")


if __name__ == "__main__":
    main(sys.argv[1], int(sys.argv[2]) if len(sys.argv) > 2 else 0, int(sys.argv[3]) if len(sys.argv) > 3 else 0)
