"""Builds libmtldas_b200.so (hand-written sm_100a CUDA + the C ABI) in-tree with nvcc."""
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
OUT = os.path.join(HERE, "mtldas_b200", "libmtldas_b200.so")
SOURCES = ["conv.cu", "elementwise.cu", "plan.cu", "api.cu", "static_conv.cu", "train_kernels.cu", "train_ew.cu", "train.cu", "stem_umma.cu", "optim_metrics.cu", "modelc.cu"]
FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17", "--use_fast_math",
         "-Xcompiler", "-fPIC,-O2,-fvisibility=hidden", "-Xptxas", "-v"]


def needs_build():
    if not os.path.exists(OUT):
        return True
    t = os.path.getmtime(OUT)
    for f in os.listdir(CSRC):
        if not os.path.isfile(os.path.join(CSRC, f)):
            continue
        if os.path.getmtime(os.path.join(CSRC, f)) > t:
            return True
    return os.path.getmtime(os.path.join(HERE, "..", "include", "mtldas.h")) > t


def build_id():
    """Short digest of every kernel source: mtldas_version() carries it, and profiles/r2_traffic.json is only trusted by
    bench.py when it was captured on the same digest."""
    import hashlib
    h = hashlib.sha1()
    for f in sorted(os.listdir(CSRC)) + [os.path.join("..", "..", "include", "mtldas.h")]:
        if not os.path.isfile(os.path.join(CSRC, f)):
            continue
        with open(os.path.join(CSRC, f), "rb") as fh:
            h.update(f.encode() + b"\0" + fh.read())
    return h.hexdigest()[:10]


def build(force=False, verbose=False):
    if not force and not needs_build():
        return OUT
    objs = []
    bid = build_id()
    os.makedirs(os.path.join(HERE, "build"), exist_ok=True)
    for s in SOURCES:
        o = os.path.join(HERE, "build", s + ".o")
        cmd = ["nvcc", *FLAGS, f'-DMTLDAS_BUILD_ID="{bid}"', "-c", os.path.join(CSRC, s), "-o", o]
        r = subprocess.run(cmd, capture_output=True, text=True)
        if verbose or r.returncode:
            sys.stderr.write(r.stdout + r.stderr)
        if r.returncode:
            raise RuntimeError("nvcc failed on " + s)
        objs.append(o)
    cmd = ["nvcc", "-shared", "-gencode", "arch=compute_100a,code=sm_100a", "-o", OUT, *objs]
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode:
        sys.stderr.write(r.stdout + r.stderr)
        raise RuntimeError("link failed")
    return OUT


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose=True))
