"""Per-CTA timeline of the tcgen05 conv kernel on a few layer shapes (MTLDAS_CONV_DEBUG=1)."""
import os, sys
os.environ["MTLDAS_CONV_DEBUG"] = "1"
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "mtl-das.pytorch_b200")]
import torch
from mtldas_b200 import conv2d_nchw
N = int(sys.argv[1]) if len(sys.argv) > 1 else 256
for (cin, cout, k, s, H, W, cin2) in [(16, 16, 3, 1, 33, 83, 0), (64, 128, 1, 2, 9, 21, 0), (128, 128, 3, 1, 5, 11, 0),
                                      (128, 64, 1, 1, 5, 11, 128), (16, 32, 3, 1, 33, 83, 0)]:
    x = torch.randn(N, cin, H, W, device="cuda")
    x2 = torch.randn(N, cin2, H, W, device="cuda") if cin2 else None
    w = torch.randn(cout, cin + cin2, k, k, device="cuda") * 0.05
    print(f"== conv {cin}+{cin2}->{cout} k{k} s{s} {H}x{W} N={N}", file=sys.stderr, flush=True)
    conv2d_nchw(x, w, None, None, x2, stride=s, epilogue=1, precision="bf16")
