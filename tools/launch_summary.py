"""Summarise an ncu launch list (--metrics gpu__time_duration.sum --csv): per-kernel totals of the LAST `frac`
of the launches (default: second half = second step of tools/one_train_step.py), optionally the ordered list."""
import collections
import csv
import re
import sys


def main():
    path = sys.argv[1]
    parts = int(sys.argv[2]) if len(sys.argv) > 2 else 2
    listing = len(sys.argv) > 3
    rows = list(csv.reader(l for l in open(path) if l.startswith('"')))
    hdr, rows = rows[0], rows[1:]
    ki, vi, gi, bi = hdr.index("Kernel Name"), hdr.index("Metric Value"), hdr.index("Grid Size"), hdr.index("Block Size")
    rows = rows[len(rows) - len(rows) // parts:]
    agg = collections.OrderedDict()
    for i, r in enumerate(rows):
        k = re.sub(r"\(.*", "", r[ki]).replace("void ", "").replace("mtldas::", "").replace("__nv_bfloat16", "bf")
        v = float(r[vi].replace(",", "")) / 1e3
        a = agg.setdefault(k, [0, 0.0])
        a[0] += 1
        a[1] += v
        if listing:
            print(i, k[:40], r[gi], r[bi], round(v, 1))
    tot = sum(a[1] for a in agg.values())
    for k, a in sorted(agg.items(), key=lambda x: -x[1][1]):
        print(f"{k[:60]:60s} n={a[0]:4d} total_us={a[1]:9.1f} avg={a[1] / a[0]:7.1f} share={a[1] / tot:.3f}")
    print("total us", round(tot, 1), "launches", len(rows))


if __name__ == "__main__":
    main()
