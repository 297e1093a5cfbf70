"""tcgen05 weight gradient vs the CUDA-core weight gradient on identical bf16 inputs (separate processes)."""
import os, subprocess, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CODE = f"""
import sys; sys.path[:0]=[{ROOT!r}, {os.path.join(ROOT, 'mtl-das.pytorch_b200')!r}]
import torch
from mtldas_b200 import TrainPlan, synth as O
n=int(sys.argv[1]); kind=sys.argv[3]
sd=O.synth_state_dict(kind,seed=0); x,yd,ye=O.synth_patches(n,seed=5)
plan=TrainPlan(kind,max(n,8),'bf16','cuda:0')
state=[v.cuda().contiguous() for v in sd.values()]
outs=plan.forward(state,x.cuda())
douts=[torch.randn_like(o)/n for o in outs] if False else [torch.full_like(o,-1.0/n) for o in outs]
flat=plan.backward(douts); torch.cuda.synchronize()
torch.save(flat.cpu(), sys.argv[2])
"""
n = sys.argv[1] if len(sys.argv) > 1 else "6"
kind = sys.argv[2] if len(sys.argv) > 2 else "MTL"
for mode, env in (("umma", {}), ("cuda", {"MTLDAS_WGRAD_CUDA_CORES": "1"})):
    r = subprocess.run([sys.executable, "-c", CODE, n, f"/tmp/wg_{mode}.pt", kind], env={**os.environ, **env}, capture_output=True, text=True, timeout=300)
    print(mode, "rc", r.returncode, r.stderr.strip().splitlines()[-3:] if r.returncode else "")
import torch
sys.path[:0] = [ROOT, os.path.join(ROOT, "mtl-das.pytorch_b200")]
from mtldas_b200 import synth as O
a, b = torch.load("/tmp/wg_umma.pt").double(), torch.load("/tmp/wg_cuda.pt").double()
print("global rel diff", float((a - b).norm() / b.norm()))
off = 0
worst = []
for k, shp, dt in O.state_dict_keys(kind):
    if k.endswith(("running_mean", "running_var", "num_batches_tracked")):
        continue
    m = 1
    for s in shp: m *= s
    da, db = a[off:off + m], b[off:off + m]; off += m
    worst.append((float((da - db).norm() / (db.norm() + 1e-20)), k, float(db.norm())))
worst.sort(reverse=True)
for w in worst[:10]: print("  %.3e %s (norm %.3e)" % w)
