"""Summarise an `ncu --metrics gpu__time_duration.sum --csv` launch list: per kernel/grid class,
launches per step, average and total device time.
usage: launch_summary.py file.csv steps [skip_steps]      (equal-length steps)
       launch_summary.py file.csv --from <kernel substring>   (the LAST forward: from the last launch of that kernel to the end)
       launch_summary.py file.csv --bench    (a `bench.py --no-extras --steps 2 --warmup 3` capture: the library launches between
                                              the last two input stages = one input + output stage and the ViT + decoder launches
                                              of two forwards, which are halved)"""
import collections
import csv
import re
import sys


def bench_slice(path):
    lines = [l for l in open(path) if not l.startswith("==")]
    seq = []
    for row in csv.DictReader(lines):
        if row.get("Metric Name") != "gpu__time_duration.sum":
            continue
        name = re.sub(r"\(.*", "", row["Kernel Name"]).replace("void ", "")
        v = float(row["Metric Value"].replace(",", ""))
        v = v / 1000 if row["Metric Unit"] == "ns" else (v * 1000 if row["Metric Unit"] == "ms" else v)
        seq.append((name, row["Grid Size"], v))
    idx = [i for i, (n, _, _) in enumerate(seq) if "input_scan" in n]
    seq = [s for s in seq[idx[-2]:idx[-1]] if s[0].startswith("mde::")]
    once = ("input_scan", "range_finalize", "patchify224", "resize_maps")
    agg = collections.OrderedDict()
    for n, g, v in seq:
        w = 1.0 if any(o in n for o in once) else 0.5
        a = agg.setdefault((n, g), [0.0, 0.0])
        a[0] += w
        a[1] += v * w
    tot = sum(t for _, t in agg.values())
    cnt = sum(c for c, _ in agg.values())
    print(f"{cnt:.0f} library launches per forward; total {tot:.1f} us/step (serialised, cold-cache)")
    print(f"source: {path} = `ncu --metrics gpu__time_duration.sum --clock-control none --launch-skip 300 -c 400 --csv python bench.py --no-extras --steps 2 --warmup 3`,")
    print("        the launches between the last two input stages (one input + output stage, the ViT + decoder launches of two forwards: halved)")
    print(f"{'us/step':>9} {'share':>6} {'n/step':>6} {'avg us':>8}  kernel grid")
    for (n, g), (c, t) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        print(f"{t:9.1f} {t / tot:6.1%} {c:6.1f} {t / c:8.2f}  {n[:60]} {g}")


def main():
    path = sys.argv[1]
    if sys.argv[2] == "--bench":
        return bench_slice(path)
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
