"""Per-launch table and per-kernel totals of an `ncu --set full` capture exported with `ncu -i X.ncu-rep --page raw --csv`:
    python tools/ncu_table.py raw.csv "<header line>" > profiles/<name>.txt"""
import collections
import csv
import re
import sys

COLS = [("gpu__time_duration.sum", "time_us"), ("sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active", "tensor%"),
        ("sm__throughput.avg.pct_of_peak_sustained_elapsed", "sm%"), ("gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "dram%"),
        ("dram__bytes_read.sum", "dram_rd_MB"), ("dram__bytes_write.sum", "dram_wr_MB"), ("lts__t_sector_hit_rate.pct", "L2hit%"),
        ("sm__warps_active.avg.pct_of_peak_sustained_active", "warps%"), ("launch__registers_per_thread", "regs")]
SCALE = {"byte": 1e-6, "Kbyte": 1e-3, "Mbyte": 1.0, "Gbyte": 1e3, "nsecond": 1e-3, "usecond": 1.0, "msecond": 1e3, "ns": 1e-3, "us": 1.0, "ms": 1e3}

rows = list(csv.reader(open(sys.argv[1])))
start = next(i for i, r in enumerate(rows) if "Kernel Name" in r)
hdr, units, data = rows[start], rows[start + 1], rows[start + 2:]
idx = {name: hdr.index(name) for name, _ in COLS if name in hdr}
ki, gi = hdr.index("Kernel Name"), hdr.index("Grid Size")
print(sys.argv[2] if len(sys.argv) > 2 else "")
print("%-4s %-34s %-14s " % ("#", "kernel", "grid") + " ".join("%10s" % short for _, short in COLS))
agg = collections.OrderedDict()
for n, r in enumerate(data):
    if len(r) <= ki:
        continue
    name = re.sub(r"\(.*", "", r[ki]).replace("void ", "").replace("mtldas::", "").replace("(anonymous namespace)::", "")
    vals = []
    for full, short in COLS:
        if full not in idx:
            vals.append(float("nan"))
            continue
        v = float(r[idx[full]].replace(",", "") or "nan")
        vals.append(v * SCALE.get(units[idx[full]], 1.0))
    print("%-4d %-34s %-14s " % (n, name[:34], r[gi].replace(" ", "")) + " ".join("%10.1f" % v for v in vals))
    a = agg.setdefault(name, [0, 0.0, 0.0, 0.0, 0.0])
    a[0] += 1
    a[1] += vals[0]
    a[2] += vals[1] * vals[0]
    a[3] += vals[4]
    a[4] += vals[5]
print()
tot = sum(a[1] for a in agg.values())
for name, a in agg.items():
    print(f"{name[:40]:40s} launches {a[0]:3d}  time {a[1]:9.1f} us ({a[1] / tot:.2f} of the capture)  tensor-pipe active (time-weighted) "
          f"{a[2] / a[1]:5.1f} %  DRAM read {a[3]:9.1f} MB  write {a[4]:9.1f} MB")
