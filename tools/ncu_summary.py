#!/usr/bin/env python
"""Summarise an `ncu --metrics gpu__time_duration.sum --csv` launch list: per-kernel count, total and share.
usage: python tools/ncu_summary.py gpurun_out/launches.csv [skip_first_n_launches]"""
import csv
import re
import sys
from collections import defaultdict


def main():
    path = sys.argv[1]
    skip = int(sys.argv[2]) if len(sys.argv) > 2 else 0
    rows = []
    with open(path, newline="") as f:
        lines = [l for l in f if not l.startswith("==")]
    rd = csv.reader(lines)
    hdr = None
    for r in rd:
        if hdr is None:
            if "Kernel Name" in r:
                hdr = r
            continue
        if len(r) != len(hdr):
            continue
        d = dict(zip(hdr, r))
        if d.get("Metric Name") != "gpu__time_duration.sum":
            continue
        v = float(d["Metric Value"].replace(",", ""))
        unit = d.get("Metric Unit", "ns")
        scale = {"ns": 1e-3, "us": 1.0, "ms": 1e3, "s": 1e6}.get(unit, 1e-3)
        rows.append((int(d["ID"]), re.sub(r"\(.*", "", d["Kernel Name"]), v * scale))
    rows = [r for r in rows if r[0] >= skip]
    agg = defaultdict(lambda: [0, 0.0])
    for _, k, us in rows:
        agg[k][0] += 1
        agg[k][1] += us
    tot = sum(v[1] for v in agg.values())
    print("launches %d, total %.1f us" % (len(rows), tot))
    for k, (n, us) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        print("%-60s n=%6d total=%12.1f us  avg=%10.1f us  share=%5.1f%%" % (k[:60], n, us, us / n, 100 * us / tot))


if __name__ == "__main__":
    main()
