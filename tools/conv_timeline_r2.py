"""Per-iteration timelines of the persistent conv kernel (MTLDAS_CONV_DEBUG) for the layers round 2 looked at.
Run under gpurun; output on stderr."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "mtl-das.pytorch_b200")]
os.environ["MTLDAS_CONV_DEBUG"] = "1"
import torch  # noqa: E402
from mtldas_b200 import conv2d_nchw  # noqa: E402

cases = [  # cin, cout, k, stride, H, W, epi, n
    (64, 64, 3, 1, 9, 21, 2, 4096),
    (128, 128, 3, 1, 5, 11, 1, 4096),
    (64, 128, 3, 1, 9, 21, 1, 4096),
    (16, 16, 3, 1, 33, 83, 1, 1024),
]
for cin, cout, k, s, H, W, epi, n in cases:
    g = torch.Generator().manual_seed(1)
    x = torch.randn(n, cin, H, W, generator=g).cuda()
    w = (torch.randn(cout, cin, k, k, generator=g) / (k * cin ** 0.5)).cuda()
    aux = torch.randn(n, cout, H, W, generator=g).cuda() if epi in (2, 3) else None
    sys.stderr.write(f"=== {cin}->{cout} k{k} s{s} {H}x{W} epi {epi} n {n} NSPLIT={os.environ.get('MTLDAS_NO_NSPLIT', '0')}\n")
    conv2d_nchw(x, w, None, aux, stride=s, epilogue=epi, precision="bf16")
