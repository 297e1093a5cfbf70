import os, sys
ROOT = "/root/repo"
sys.path[:0] = [ROOT, os.path.join(ROOT, "mtl-das.pytorch_b200")]
os.environ["MTLDAS_CONV_DEBUG"] = "1"
import torch
from mtldas_b200 import conv2d_nchw
cases = [(16, 32, 3, 1, 33, 83, 5, 2048), (16, 32, 3, 1, 33, 83, 1, 2048), (8, 16, 3, 1, 33, 83, 3, 2048), (16, 16, 3, 1, 33, 83, 2, 2048), (32, 64, 3, 1, 17, 42, 5, 4096)]
for cin, cout, k, s, H, W, epi, n in cases:
    g = torch.Generator().manual_seed(1)
    x = torch.randn(n, cin, H, W, generator=g).cuda()
    w = (torch.randn(cout, cin, k, k, generator=g) / (k * cin ** 0.5)).cuda()
    aux = torch.randn(n, cout, H, W, generator=g).cuda() if epi in (2, 3) else None
    sys.stderr.write(f"=== {cin}->{cout} k{k} s{s} {H}x{W} epi {epi} n {n}\n")
    conv2d_nchw(x, w, None, aux, stride=s, epilogue=epi, precision="bf16")
    torch.cuda.synchronize()
