"""Summarise an `ncu --metrics gpu__time_duration.sum --csv` launch list: per-kernel count,
total device time and share.  usage: python profiles/tools_launch_summary.py launches.csv"""
import collections
import csv
import sys


def main(path):
    with open(path) as f:
        lines = [l for l in f if not l.startswith("==")]
    agg = collections.defaultdict(lambda: [0, 0.0])
    for row in csv.DictReader(lines):
        if row.get("Metric Name") != "gpu__time_duration.sum":
            continue
        v = float(row["Metric Value"].replace(",", ""))
        u = row["Metric Unit"]
        v = v / 1000 if u == "ns" else v * 1000 if u == "ms" else v
        k = row["Kernel Name"][:110]
        agg[k][0] += 1
        agg[k][1] += v
    tot = sum(v[1] for v in agg.values())
    print(f"total {tot:.1f} us over {sum(v[0] for v in agg.values())} launches")
    for k, v in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        print(f"{v[1]:10.1f} us {100 * v[1] / tot:5.1f}%  x{v[0]:5d}  avg {v[1] / v[0]:7.2f} us  {k}")


if __name__ == "__main__":
    main(sys.argv[1])
