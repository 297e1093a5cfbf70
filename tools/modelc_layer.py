"""Time ONE model C convolution through the C ABI: python tools/modelc_layer.py n cin cout k h w [stride pad]"""
import ctypes
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "mtl-das.pytorch_b200")]
import torch  # noqa: E402
from mtldas_b200 import lib  # noqa: E402

n, cin, cout, kh, kw, h, w = (int(v) for v in sys.argv[1:8])
stride = int(sys.argv[8]) if len(sys.argv) > 8 else 1
ph, pw = (int(sys.argv[9]), int(sys.argv[10])) if len(sys.argv) > 10 else (0, 0)
ho, wo = (h + 2 * ph - kh) // stride + 1, (w + 2 * pw - kw) // stride + 1
x = torch.randn(cin // 8, n * h * w, 8, device="cuda").bfloat16()
wp = torch.randn(kh * kw, cin // 8, cout, 8, device="cuda").bfloat16()
b = torch.randn(cout, device="cuda")
out = torch.empty(cout // 8, n * ho * wo, 8, device="cuda", dtype=torch.bfloat16)
st = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
P = lambda t: ctypes.c_void_p(t.data_ptr())  # noqa: E731
reps = int(os.environ.get("REPS", "5"))
for it in range(2):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        lib.mtldas_c_conv(P(x), n, cin, h, w, P(wp), P(b), P(out), cout, 0, cout, kh, kw, stride, stride, ph, pw, 1, st)
    e1.record()
    torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / reps
print(f"{cin}->{cout} k{kh}x{kw} s{stride} {h}x{w} n={n}: {ms:.4f} ms  {2.0 * n * ho * wo * cout * cin * kh * kw / ms / 1e9:.1f} TFLOP/s")
