"""Training forward/backward against the fp64 oracle over awkward batch sizes (ragged tiles, tiny grids)."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "mtl-das.pytorch_b200")]
import torch
from oracle import mtl_oracle as O
from mtldas_b200 import TrainPlan, nll_loss

sd = O.synth_state_dict("MTL", seed=0)
names = O.param_names("MTL")
for precision, sizes in (("bf16", (1, 2, 5, 32, 33, 100)), ("fp32", (1, 3, 32))):
    for n in sizes:
        x, yd, ye = O.synth_patches(n, seed=100 + n)
        _, rloss, rgrads, _ = O.loss_and_grads(sd, x, yd, ye, "MTL", dtype=torch.float64)
        ref = torch.cat([rgrads[k].reshape(-1) for k in names]).double()
        plan = TrainPlan("MTL", n, precision, "cuda:0")
        state = [v.cuda().contiguous() for v in sd.values()]
        res = []
        for rep in range(3):   # eager, captured, replayed
            outs = plan.forward(state, x.cuda())
            loss, douts = nll_loss(list(outs), [yd.cuda(), ye.cuda()])
            flat = plan.backward(douts).cpu().double()
            res.append((float(loss), flat))
        cos = float(torch.dot(res[0][1], ref) / (res[0][1].norm() * ref.norm()))
        rel = float((res[0][1] - ref).norm() / ref.norm())
        same = all(torch.equal(res[0][1] != res[0][1], r[1] != r[1]) for r in res) and max(abs(r[0] - res[0][0]) for r in res) < 1e-5
        print(f"{precision} n={n:4d} loss {res[0][0]:.5f} (oracle {float(rloss):.5f}) grad cos {cos:.5f} rel-L2 {rel:.2e} "
              f"nan {bool(torch.isnan(res[0][1]).any())} repeatable {same}", flush=True)
