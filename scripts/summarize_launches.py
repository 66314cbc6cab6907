#!/usr/bin/env python
"""Aggregate an `ncu --metrics gpu__time_duration.sum --csv` launch list per kernel name.

    python scripts/summarize_launches.py gpurun_out/launches.csv > profiles/rNN_launches_summary.txt
"""
import collections
import csv
import re
import sys


def main(path):
    with open(path) as f:
        lines = [l for l in f if not l.startswith("==")]
    agg = collections.OrderedDict()
    for row in csv.DictReader(lines):
        if row.get("Metric Name") != "gpu__time_duration.sum":
            continue
        name = re.sub(r"\(.*", "", row["Kernel Name"])
        name = name.replace("void ", "").replace("b200spai::", "")
        ours = not any(t in row["Kernel Name"] for t in ("cub::", "at::", "thrust::", "nccl"))
        v = float(row["Metric Value"].replace(",", ""))
        v *= {"ns": 1e-6, "us": 1e-3, "ms": 1.0, "s": 1e3}.get(row["Metric Unit"], 1.0)
        a = agg.setdefault(name[:90], [0, 0.0, ours])
        a[0] += 1
        a[1] += v
    tot = sum(v[1] for v in agg.values())
    ours = sum(v[1] for v in agg.values() if v[2])
    print(f"# {path}: {sum(v[0] for v in agg.values())} launches, {tot:.3f} ms summed device time "
          f"(cold-cache, serialised); libb200spai kernels {ours:.3f} ms = {100 * ours / max(tot, 1e-9):.1f} %")
    print(f"{'kernel':92s} {'launches':>8s} {'ms':>10s} {'share':>7s} {'us/launch':>10s}")
    for k, (c, t, o) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        print(f"{k:92s} {c:8d} {t:10.3f} {100 * t / tot:6.1f}% {1e3 * t / c:10.2f}{'' if o else '   (torch/cub: input generation + sort/scan)'}")


if __name__ == "__main__":
    main(sys.argv[1])
