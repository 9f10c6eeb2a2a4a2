#!/usr/bin/env python
"""Summarises an `ncu --metrics gpu__time_duration.sum --csv` launch list per kernel (total ms, share, launches).
usage: python tools/launch_summary.py gpurun_out/launches.csv "<command that was profiled>" > profiles/<name>.txt"""
import csv
import re
import sys
from collections import defaultdict


def main():
    path, cmd = sys.argv[1], (sys.argv[2] if len(sys.argv) > 2 else "?")
    rows = [l for l in open(path) if l.startswith('"')]
    tot, cnt = defaultdict(float), defaultdict(int)
    for r in csv.DictReader(rows):
        if r["Metric Name"] != "gpu__time_duration.sum":
            continue
        ns = float(r["Metric Value"].replace(",", ""))
        if r["Metric Unit"] in ("us", "usecond"):
            ns *= 1e3
        elif r["Metric Unit"] in ("ms", "msecond"):
            ns *= 1e6
        name = re.sub(r"\(.*$", "", r["Kernel Name"])[:70]
        tot[name] += ns / 1e6
        cnt[name] += 1
    total = sum(tot.values())
    print(f"{sum(cnt.values())} launches of `{cmd}` under ncu --metrics gpu__time_duration.sum --clock-control none; total {total:.1f} ms")
    print("per-launch times are cold-cache and serialised: compare SHARES with bench.py's stage timers, not absolutes\n")
    print("  total ms  share launches  ms/launch  kernel")
    for k in sorted(tot, key=lambda k: -tot[k]):
        print(f"{tot[k]:10.3f} {100 * tot[k] / total:5.1f}% {cnt[k]:8d} {tot[k] / cnt[k]:10.4f}  {k}")


if __name__ == "__main__":
    main()
