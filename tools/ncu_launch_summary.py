#!/usr/bin/env python
"""Summarise an `ncu --metrics gpu__time_duration.sum --csv` launch list:
launches, total device time and share per kernel (functor) name.

    python tools/ncu_launch_summary.py gpurun_out/launches.csv
"""
import collections
import csv
import re
import sys


def main(path):
    with open(path) as f:
        lines = [ln for ln in f if not ln.startswith("==")]
    r = csv.reader(lines)
    hdr = next(r)
    ki, vi, ui = (hdr.index(c) for c in
                  ("Kernel Name", "Metric Value", "Metric Unit"))
    agg = collections.defaultdict(lambda: [0, 0.0])
    for row in r:
        if len(row) <= vi:
            continue
        v = float(row[vi].replace(",", ""))
        v *= {"ns": 1e-3, "us": 1.0, "ms": 1e3, "s": 1e6}.get(row[ui], 1.0)
        name = row[ki]
        # "void k_for_each<unnamed>::OrderKernel>(T1, unsigned long)"
        short = name.split(">(")[0].split("(")[0]
        short = re.sub(r"^void ", "", short)
        short = short.replace("<unnamed>::", "").replace(
            "(anonymous namespace)::", "").replace("sanm_b200::", "")
        agg[short][0] += 1
        agg[short][1] += v
    tot = sum(v[1] for v in agg.values())
    n = sum(v[0] for v in agg.values())
    print(f"# {path}: {n} launches, {tot / 1e3:.3f} ms of kernel time "
          "(ncu per-launch times: serialised, cold cache)")
    print(f"{'kernel':44s} {'launches':>8s} {'ms':>10s} {'share':>7s} "
          f"{'avg_us':>9s}")
    for k, v in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        print(f"{k:44s} {v[0]:8d} {v[1] / 1e3:10.3f} {v[1] / tot:7.3f} "
              f"{v[1] / v[0]:9.1f}")


if __name__ == "__main__":
    main(sys.argv[1])
