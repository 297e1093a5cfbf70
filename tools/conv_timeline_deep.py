import os, sys
os.environ["MTLDAS_CONV_DEBUG"] = "1"
sys.path[:0] = ["/root/repo", "/root/repo/mtl-das.pytorch_b200"]
import torch
from mtldas_b200 import conv2d_nchw
N = 1024
for (cin, cout, k, s, H, W, cin2) in [(128, 128, 3, 1, 5, 11, 0), (64, 128, 3, 1, 9, 21, 0), (64, 64, 3, 1, 9, 21, 0)]:
    x = torch.randn(N, cin, H, W, device="cuda")
    w = torch.randn(cout, cin, k, k, device="cuda") * 0.05
    print(f"== conv {cin}->{cout} k{k} s{s} {H}x{W} N={N}", file=sys.stderr, flush=True)
    conv2d_nchw(x, w, None, None, None, stride=s, epilogue=1, precision="bf16")
