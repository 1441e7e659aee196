"""Summarise an `ncu --metrics gpu__time_duration.sum --csv` launch list: per kernel/grid class,
launches per step, average and total device time.
usage: launch_summary.py file.csv steps [skip_steps]      (equal-length steps)
       launch_summary.py file.csv --from <kernel substring>   (the LAST forward: from the last launch of that kernel to the end)"""
import collections
import csv
import re
import sys


def main():
    path = sys.argv[1]
    start_name = sys.argv[3] if sys.argv[2] == "--from" else None
    steps = 1 if start_name else int(sys.argv[2])
    skip = int(sys.argv[3]) if len(sys.argv) > 3 and not start_name else 0
    lines = [l for l in open(path) if not l.startswith("==")]
    seq = []
    for row in csv.DictReader(lines):
        if row.get("Metric Name") != "gpu__time_duration.sum":
            continue
        name = re.sub(r"\(.*", "", row["Kernel Name"]).replace("void ", "")
        v = float(row["Metric Value"].replace(",", ""))
        v = v / 1000 if row["Metric Unit"] == "ns" else (v * 1000 if row["Metric Unit"] == "ms" else v)
        seq.append((name, row["Grid Size"], v))
    if start_name:
        first = max(i for i, (n, _, _) in enumerate(seq) if start_name in n)
        seq = seq[first:]
        per = len(seq)
    else:
        per = len(seq) // (steps + skip)
        seq = seq[per * skip:per * (skip + steps)]
    agg = collections.OrderedDict()
    for n, g, v in seq:
        a = agg.setdefault((n, g), [0, 0.0])
        a[0] += 1
        a[1] += v
    tot = sum(v for _, _, v in seq)
    print(f"{len(seq)} launches over {steps} steps ({per}/step); total {tot / steps:.1f} us/step (serialised, cold-cache)")
    print(f"{'us/step':>9} {'share':>6} {'n/step':>6} {'avg us':>8}  kernel grid")
    for (n, g), (c, t) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        print(f"{t / steps:9.1f} {t / tot:6.1%} {c / steps:6.1f} {t / c:8.2f}  {n[:60]} {g}")


if __name__ == "__main__":
    main()
