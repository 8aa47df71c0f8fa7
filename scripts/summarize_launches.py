#!/usr/bin/env python3
"""Summarise an `ncu --metrics gpu__time_duration.sum --csv` launch list per kernel (count, total, share)."""
import collections
import csv
import sys


def main(path):
    rows = list(csv.reader(open(path)))
    hdr = None
    agg = collections.defaultdict(lambda: [0, 0.0])
    for r in rows:
        if "Kernel Name" in r:
            hdr = r
            continue
        if hdr and len(r) == len(hdr):
            d = dict(zip(hdr, r))
            if d.get("Metric Name") == "gpu__time_duration.sum":
                v = float(d["Metric Value"].replace(",", ""))
                if d["Metric Unit"] in ("ns", "nsecond"):
                    v /= 1e3
                k = d["Kernel Name"].split("(")[0][:70]
                agg[k][0] += 1
                agg[k][1] += v
    tot = sum(v[1] for v in agg.values())
    print("# %s  (per-launch times are cold-cache and serialised under ncu: compare SHARES)" % path)
    for k, v in sorted(agg.items(), key=lambda x: -x[1][1]):
        print("%-72s n=%5d total %12.1f us  avg %10.1f us  share %.4f" % (k, v[0], v[1], v[1] / v[0], v[1] / tot))


if __name__ == "__main__":
    main(sys.argv[1])
