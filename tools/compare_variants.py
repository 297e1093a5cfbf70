"""Table of the bench variants written by tools/bench_variants.sh (per-op milliseconds side by side)."""
import glob
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
base = json.loads(open(os.path.join(ROOT, sys.argv[1] if len(sys.argv) > 1 else "gpurun_out/r2_bench_1.json")).read().strip().splitlines()[-1])
runs = [("base", base)]
for f in sorted(glob.glob(os.path.join(ROOT, "gpurun_out", "variant_*.json"))):
    try:
        runs.append((open(f.replace(".json", ".env")).read().strip(), json.loads(open(f).read().strip().splitlines()[-1])))
    except Exception as e:
        print(f, "unreadable", e, open(f.replace(".json", ".err")).read()[-400:])
for i, (n, x) in enumerate(runs):
    print(i, n, "| samples/s", round(x["value"]), "ms", round(x["ms_per_step"], 3), "frac", round(x["roofline"]["frac"], 3),
          "e2e", round(x["e2e"]["value"]), "bf16up", round((x.get("e2e_bf16_upload") or {}).get("value", 0)))
tab = {}
for i, (n, x) in enumerate(runs):
    seen = {}
    for k, v in x["roofline"]["per_op_ms"].items():
        name = k.rsplit("#", 1)[0]
        seen[name] = seen.get(name, 0) + 1
        tab.setdefault(name + ("" if seen[name] == 1 else "'"), {})[i] = v
for k, row in tab.items():
    print(f"{k:26s}", "  ".join(f"{row.get(i, float('nan')):7.4f}" for i in range(len(runs))))
