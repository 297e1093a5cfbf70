"""Summarise an `ncu --set full` capture of the conv launches of one eager MTL forward (tools/one_forward.py).

    ncu -i gpurun_out/r2_prof.ncu-rep --page raw --csv > gpurun_out/r2_prof_raw.csv      (here, no GPU needed)
    python tools/ncu_traffic_summary.py gpurun_out/r2_prof_raw.csv 4096 bf16 "<mtldas_version of the profiled build>"

writes profiles/r2_conv_ncu_full_batch<N>.txt (per-launch table) and profiles/r2_traffic.json, which bench.py reads for
`roofline.traffic` -- only when its `mtldas_version` equals the running library's (a stale profile is never quoted)."""
import csv
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
COLS = ["Kernel Name", "Grid Size", "Block Size", "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active",
        "launch__registers_per_thread", "lts__t_sector_hit_rate.pct", "sm__inst_executed.avg.per_cycle_elapsed"]


def to_bytes(v, unit):
    v = float(v.replace(",", ""))
    return v * {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}.get(unit, 1)


def to_us(v, unit):
    v = float(v.replace(",", ""))
    return v * {"ns": 1e-3, "us": 1, "ms": 1e3, "s": 1e6, "nsecond": 1e-3, "usecond": 1, "msecond": 1e3, "second": 1e6}.get(unit, 1)


def main():
    path, batch, precision, version = sys.argv[1], int(sys.argv[2]), sys.argv[3], sys.argv[4]
    rows = list(csv.reader(open(path)))
    while rows and "Kernel Name" not in rows[0]:
        rows.pop(0)
    head, units, body = rows[0], rows[1], rows[2:]
    idx = {c: head.index(c) for c in COLS if c in head}
    out, tot_r, tot_w, tot_t, n = [], 0.0, 0.0, 0.0, 0
    for r in body:
        if len(r) < len(head) or "conv_umma" not in r[idx["Kernel Name"]]:
            continue
        rd = to_bytes(r[idx["dram__bytes_read.sum"]], units[idx["dram__bytes_read.sum"]])
        wr = to_bytes(r[idx["dram__bytes_write.sum"]], units[idx["dram__bytes_write.sum"]])
        t = to_us(r[idx["gpu__time_duration.sum"]], units[idx["gpu__time_duration.sum"]])
        tot_r, tot_w, tot_t, n = tot_r + rd, tot_w + wr, tot_t + t, n + 1
        cells = [r[idx["Kernel Name"]].split("(")[0], r[idx["Grid Size"]], r[idx["Block Size"]], f"{t:.1f}", f"{rd / 1e6:.1f}", f"{wr / 1e6:.1f}"]
        cells += [r[idx[c]] for c in COLS[6:] if c in idx]
        out.append(" | ".join(cells))
    txt = os.path.join(ROOT, "profiles", f"r2_conv_ncu_full_batch{batch}.txt")
    with open(txt, "w") as f:
        f.write("MTLDAS_NO_GRAPH=1 ncu --set full --clock-control none -k regex:conv_umma --launch-skip 30 -c 30 "
                f"python tools/one_forward.py {batch}\n(the conv launches of ONE eager MTL forward at batch {batch}, {precision}, build {version}; "
                "per-launch values, cold cache, serialised)\n")
        f.write(f"sum over the {n} conv launches: dram read {tot_r / 1e6:.0f} MB + write {tot_w / 1e6:.0f} MB = {(tot_r + tot_w) / 1e6:.0f} MB; time {tot_t:.0f} us\n")
        f.write("kernel | grid | block | time(us) | dram read(MB) | dram write(MB) | " + " | ".join(c for c in COLS[6:] if c in idx) + "\n")
        f.write("\n".join(out) + "\n")
    js = {"mtldas_version": version, "batch": batch, "precision": precision, "conv_launches_per_step": n,
          "conv_dram_bytes_per_step": tot_r + tot_w, "conv_time_us_under_ncu": tot_t, "source": os.path.relpath(txt, ROOT)}
    json.dump(js, open(os.path.join(ROOT, "profiles", "r2_traffic.json"), "w"), indent=1)
    print(open(txt).read())


if __name__ == "__main__":
    main()
