"""Compare training forward/backward between the tensor-core and the CUDA-core stem (two subprocesses)."""
import os, subprocess, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CODE = r'''
import sys, os, torch
sys.path[:0] = [%r, %r]
from mtldas_b200 import TrainPlan, synth as O
n = 16
sd = O.synth_state_dict("MTL", seed=0, trained_like=False)
x, yd, ye = O.synth_patches(n, seed=500)
plan = TrainPlan("MTL", n, "bf16", "cuda:0")
state = [v.cuda().contiguous() for v in sd.values()]
xs = x.cuda()
d0 = torch.zeros(n, 16, device="cuda"); d0[torch.arange(n), yd.cuda()] = -1.0 / n
d1 = torch.zeros(n, 2, device="cuda"); d1[torch.arange(n), ye.cuda()] = -1.0 / n
flat = torch.empty(plan.n_params, device="cuda")
res = []
for it in range(3):
    outs = plan.forward(state, xs)
    plan.backward([d0, d1], flat)
    torch.cuda.synchronize()
    res.append((outs[0].cpu().clone(), outs[1].cpu().clone(), flat.cpu().clone()))
torch.save(res, sys.argv[1])
''' % (ROOT, os.path.join(ROOT, "mtl-das.pytorch_b200"))
import torch
for flag in ("1", "0"):
    env = dict(os.environ, MTLDAS_UMMA_STEM=flag)
    subprocess.run([sys.executable, "-c", CODE, f"/tmp/stemchk{flag}.pt"], env=env, check=True, timeout=300)
a, b = torch.load("/tmp/stemchk1.pt"), torch.load("/tmp/stemchk0.pt")
for it in range(3):
    o = [float((a[it][k] - b[it][k]).abs().max()) for k in range(2)]
    g = a[it][2] - b[it][2]
    print(f"iter {it}: out diff {o}  grad rel-L2 {float(g.norm() / b[it][2].norm()):.3e}  first-784 (stem w) rel "
          f"{float(g[:784].norm() / b[it][2][:784].norm()):.3e}  nan {bool(torch.isnan(a[it][2]).any())}")
for name, r in (("umma", a), ("ffma", b)):
    print(name, "self-consistency out diff iter1-iter0", float((r[1][0] - r[0][0]).abs().max()), "iter2-iter0",
          float((r[2][0] - r[0][0]).abs().max()))
