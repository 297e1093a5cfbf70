import os, sys, time
ROOT = "/root/repo"
sys.path[:0] = [ROOT, os.path.join(ROOT, "mtl-das.pytorch_b200")]
import torch
from mtldas_b200 import MTL_Net, synth as O, nll_loss
from mtldas_b200.ddp import DataParallelTrainer
nb = 256
dev = torch.device("cuda:0")
m = MTL_Net(); m.load_state_dict(O.synth_state_dict("MTL", seed=0)); m = m.to(dev).train()
opt = torch.optim.Adam(m.parameters(), lr=1e-3 / 1.5, weight_decay=1e-5, fused=True)
tr = DataParallelTrainer(m, opt)
x, yd, ye = O.synth_patches(64, seed=1)
x = x.repeat(4, 1, 1, 1).to(dev); yd = yd.repeat(4).to(dev); ye = ye.repeat(4).to(dev)
def run(tag, fn, k=10):
    for _ in range(4): fn()
    torch.cuda.synchronize()
    e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
    t0 = time.perf_counter(); e0.record()
    for _ in range(k): fn()
    e1.record(); t1 = time.perf_counter(); torch.cuda.synchronize()
    print(tag, "device ms/step", round(e0.elapsed_time(e1) / k, 3), "host enqueue ms/step", round((t1 - t0) * 1e3 / k, 3), flush=True)
run("trainer.step", lambda: tr.step(x, [yd, ye]))
plan = m._train_plan(dev, nb)
state = [v.detach() for v in m.state_dict(keep_vars=True).values()]
flat = torch.empty(plan.n_params, device=dev)
bufs = [None]
def core():
    outs = plan.forward(state, x)
    bufs[0] = nll_loss(list(outs), [yd, ye], out=bufs[0])
    plan.backward(bufs[0][1], flat)
run("fwd+loss+bwd", core)
def core2():
    outs = plan.forward(state, x)
    plan.backward(bufs[0][1], flat)
run("fwd+bwd", core2)
run("state_dict()", lambda: [v.detach() for v in m.state_dict(keep_vars=True).values()])
run("adam", lambda: opt.step())
