import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "mtl-das.pytorch_b200")]
import torch
from mtldas_b200 import TrainPlan, synth as O
n = int(sys.argv[1]) if len(sys.argv) > 1 else 256
for precision in ("bf16", "fp32"):
    sd = O.synth_state_dict("MTL", seed=0)
    x, yd, ye = O.synth_patches(min(n, 64), seed=1)
    x = x.repeat((n + 63) // 64, 1, 1, 1)[:n].cuda()
    plan = TrainPlan("MTL", n, precision, "cuda:0")
    state = [v.cuda().contiguous() for v in sd.values()]
    d0 = torch.full((n, 16), -1.0 / n, device="cuda"); d1 = torch.full((n, 2), -1.0 / n, device="cuda")
    flat = torch.empty(plan.n_params, device="cuda")
    for _ in range(3):
        plan.forward(state, x); plan.backward([d0, d1], flat)
    torch.cuda.synchronize()
    e = [torch.cuda.Event(enable_timing=True) for _ in range(3)]
    K = 5
    tf = tb = 0.0
    for _ in range(K):
        e[0].record(); plan.forward(state, x); e[1].record(); plan.backward([d0, d1], flat); e[2].record()
        torch.cuda.synchronize()
        tf += e[0].elapsed_time(e[1]); tb += e[1].elapsed_time(e[2])
    print(f"{precision} batch {n}: fwd {tf/K:.2f} ms bwd {tb/K:.2f} ms -> {n/((tf+tb)/K)*1e3:.0f} samples/s; ws {plan.workspace_bytes/1e9:.2f} GB")
