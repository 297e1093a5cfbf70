"""Per-kernel counts of the SASS mnemonics that prove a Blackwell-native kernel (B200_PROFILING.md): UTCHMMA
(tcgen05.mma), LDTM (tcgen05.ld), UBLKCP (cp.async.bulk), UTCBAR (tcgen05.commit), LDGSTS (cp.async), HMMA (legacy).
    python tools/sass_summary.py > profiles/r2_sass_summary.txt"""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
so = os.path.join(ROOT, "mtl-das.pytorch_b200", "mtldas_b200", "libmtldas_b200.so")
sass = subprocess.run(["cuobjdump", "-sass", so], capture_output=True, text=True).stdout
ops = ("UTCHMMA", "LDTM", "UBLKCP", "UTCBAR", "LDGSTS", "HMMA.", "SYNCS")
counts, cur = collections.OrderedDict(), None
for line in sass.splitlines():
    m = re.search(r"Function : (\S+)", line)
    if m:
        cur = subprocess.run(["c++filt", m.group(1)], capture_output=True, text=True).stdout.strip().replace("(anonymous namespace)::", "").split("(")[0]
        counts[cur] = collections.Counter()
        continue
    if cur:
        for o in ops:
            if re.search(r"\b" + re.escape(o), line):
                counts[cur][o] += 1
sys.path.insert(0, os.path.join(ROOT, "mtl-das.pytorch_b200"))
print("cuobjdump -sass mtl-das.pytorch_b200/mtldas_b200/libmtldas_b200.so | per-kernel mnemonic counts (sm_100a)")
print(f"{'kernel':78s} " + " ".join(f"{o.rstrip('.'):>8s}" for o in ops))
tot = collections.Counter()
for k, c in counts.items():
    tot.update(c)
    if sum(c.values()):
        print(f"{k[:78]:78s} " + " ".join(f"{c[o]:8d}" for o in ops))
print(f"{'TOTAL':78s} " + " ".join(f"{tot[o]:8d}" for o in ops))
