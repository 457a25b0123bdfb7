#!/usr/bin/env python
"""Summarise an `ncu --metrics gpu__time_duration.sum --csv` launch list per kernel (count, total, median, share).

  python tools/launch_summary.py gpurun_out/launches.csv "comment line" > profiles/rNN_xxx_launches.csv
"""
import collections
import csv
import sys


def main():
    rows = list(csv.reader(open(sys.argv[1])))
    hdr, agg = None, collections.OrderedDict()
    for r in rows:
        if "Kernel Name" in r:
            hdr = r
            continue
        if hdr is None or len(r) != len(hdr):
            continue
        d = dict(zip(hdr, r))
        if d.get("Metric Name") != "gpu__time_duration.sum":
            continue
        scale = {"ns": 1e-3, "nsecond": 1e-3, "us": 1.0, "usecond": 1.0, "ms": 1e3, "msecond": 1e3}.get(d["Metric Unit"], 1e-3)
        agg.setdefault(d["Kernel Name"].split("(")[0][:90], []).append(float(d["Metric Value"].replace(",", "")) * scale)
    tot = sum(sum(v) for v in agg.values())
    for c in sys.argv[2:]:
        print("# " + c)
    print(f"# total {tot:.1f} us over {sum(len(v) for v in agg.values())} launches (cold-cache, serialised: compare SHARES)")
    print("kernel,launches,total_us,median_us,share")
    for k, v in sorted(agg.items(), key=lambda kv: -sum(kv[1])):
        print(f"\"{k}\",{len(v)},{sum(v):.1f},{sorted(v)[len(v) // 2]:.1f},{sum(v) / tot:.3f}")


if __name__ == "__main__":
    main()
