import os, sys
ROOT = "/root/repo"
sys.path[:0] = [ROOT, os.path.join(ROOT, "mtl-das.pytorch_b200")]
import torch
from mtldas_b200 import MTL_Net, synth as O
m = MTL_Net(); m.load_state_dict(O.synth_state_dict("MTL", seed=0)); m = m.cuda().eval()
base, _, _ = O.synth_patches(64, seed=3)
x = base.repeat(65, 1, 1, 1)[:4100].cuda()
x += torch.randn_like(x) * 0.1
with torch.no_grad():
    a = [o.clone() for o in m(x)]
    b1 = [o.clone() for o in m(x[:1])]
    b2 = [o.clone() for o in m(x[1:1000])]
    b3 = [o.clone() for o in m(x[1000:4100])]
for t in range(2):
    cat = torch.cat([b1[t], b2[t], b3[t]])
    print("task", t, a[t].shape, "max diff vs sliced runs", float((a[t] - cat).abs().max()), "nan", bool(torch.isnan(a[t]).any()))
