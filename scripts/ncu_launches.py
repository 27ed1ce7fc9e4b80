#!/usr/bin/env python
"""Aggregate an `ncu --metrics gpu__time_duration.sum --csv` launch list per kernel: launches, total
device time and SHARE of the captured window (per-launch times are cold-cache and serialised: compare
shares, not absolutes).   python scripts/ncu_launches.py gpurun_out/launches.csv"""
import csv
import re
import sys
from collections import defaultdict


def main():
    rows = [r for r in csv.reader(open(sys.argv[1], errors="ignore")) if len(r) > 10]
    hdr = rows[0]
    ik, iv, iu = hdr.index("Kernel Name"), hdr.index("Metric Value"), hdr.index("Metric Unit")
    tot, cnt = defaultdict(float), defaultdict(int)
    for r in rows[1:]:
        name = re.sub(r"\(.*", "", r[ik]).replace("void ", "").replace("mcpt::", "")
        ns = float(r[iv].replace(",", "")) * {"ns": 1, "us": 1e3, "usecond": 1e3, "ms": 1e6}.get(r[iu], 1)
        tot[name] += ns
        cnt[name] += 1
    total = sum(tot.values())
    print(f"# {sys.argv[1]}: {sum(cnt.values())} launches, {total / 1e6:.3f} ms of device time in the captured window")
    print(f"{'kernel':70s} {'launches':>8s} {'total ms':>10s} {'avg us':>9s} {'share':>7s}")
    for k in sorted(tot, key=tot.get, reverse=True):
        print(f"{k[:70]:70s} {cnt[k]:8d} {tot[k] / 1e6:10.3f} {tot[k] / cnt[k] / 1e3:9.2f} {100 * tot[k] / total:6.2f}%")


if __name__ == "__main__":
    main()
