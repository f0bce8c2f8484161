#!/usr/bin/env python
"""Developer tool: per-kernel totals of an ncu launch list (`ncu --metrics gpu__time_duration.sum --csv`).
  python tools/launch_summary.py gpurun_out/launches.csv > profiles/xxx_launches_summary.txt"""
import csv
import sys
from collections import defaultdict


def main():
    rows = [r for r in csv.reader(open(sys.argv[1], newline="")) if len(r) > 14 and r[0].isdigit()]
    tot, cnt = defaultdict(float), defaultdict(int)
    for r in rows:
        name = r[4].replace("void ", "").replace("zg::<unnamed>::", "").split("(")[0]
        tot[name] += float(r[14])
        cnt[name] += 1
    allt = sum(tot.values()) or 1.0
    print(f"{len(rows)} launches, {allt/1e6:.3f} ms of kernel time (cold-cache, serialised by ncu: shares matter, not absolutes)")
    for k in sorted(tot, key=lambda k: -tot[k]):
        print(f"{tot[k]/allt*100:6.2f} %  {tot[k]/1e6:10.3f} ms  {cnt[k]:5d} x  {k}")


if __name__ == "__main__":
    main()
