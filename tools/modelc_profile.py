"""Per-op device time of the model C layer program (CUDA events around every launch), with the achieved TFLOP/s of each
convolution: python tools/modelc_profile.py [batch]"""
import ctypes
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "mtl-das.pytorch_b200")]
import torch  # noqa: E402
from mtldas_b200 import lib, synth  # noqa: E402
from mtldas_b200.nets_c import Multi_Classifier, _SPECS, _out_hw  # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 1024
m = Multi_Classifier(init_weights=False)
m.load_state_dict(synth.synth_state_dict_c(0), strict=True)
m = m.cuda().eval()
x = synth.synth_patches(64, seed=1)[0].cuda().repeat((n + 63) // 64, 1, 1, 1)[:n].contiguous()
for _ in range(2):
    m(x)
torch.cuda.synchronize()
prog, packed = m._last_program, m._pack(x.device)
st = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
ptr = lambda name: ctypes.c_void_p(x.data_ptr() if name == "x" else prog.buf[name].data_ptr())  # noqa: E731
rows = []
for kind, src, dst, c0, arg in prog.ops:
    c, h, w = prog.shapes[src]
    ctot = prog.shapes[dst][0]
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(3):
        if kind == "conv":
            cin, cout, k, s, pad = _SPECS[arg]
            wp, bp, _ = packed[arg]
            lib.mtldas_c_conv(ptr(src), n, cin, h, w, ctypes.c_void_p(wp.data_ptr()), ctypes.c_void_p(bp.data_ptr()), ptr(dst),
                              ctot, c0, cout, k[0], k[1], s[0], s[1], pad[0], pad[1], 1, st)
        else:
            stride, pad, avg = arg
            lib.mtldas_c_pool(ptr(src), n, c, h, w, ptr(dst), ctot, c0, stride, pad, avg, st)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 3
    if kind == "conv":
        cin, cout, k, s, pad = _SPECS[arg]
        ho, wo = _out_hw(h, w, k, s, pad)
        flop = 2.0 * n * ho * wo * cout * cin * k[0] * k[1]
        rows.append((ms, arg, f"{cin}->{cout} k{k[0]}x{k[1]} s{s[0]} {h}x{w}", flop / ms / 1e9))
    else:
        rows.append((ms, "pool:" + dst, f"{c} ch {h}x{w}", 0.0))
tot = sum(r[0] for r in rows)
print(f"batch {n}: sum of per-op times {tot:.3f} ms -> {n / tot * 1e3:.0f} patches/s (serialised)")
for ms, name, what, tf in sorted(rows, reverse=True)[:int(os.environ.get("TOP", "40"))]:
    print(f"{ms:8.4f} ms {100 * ms / tot:5.1f}%  {tf:7.1f} TFLOP/s  {name:32s} {what}")
by = {}
for ms, name, what, tf in rows:
    key = name.split(".")[0] if not name.startswith("pool") else "pools"
    by[key] = by.get(key, 0) + ms
print({k: round(v, 3) for k, v in by.items()})
