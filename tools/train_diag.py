"""Training-path diagnostics on a GPU box: prints errors of forward outputs, running statistics, saved features,
feature gradients and parameter gradients against the fp64 oracle (no asserts)."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "mtl-das.pytorch_b200")]
import torch
from oracle import mtl_oracle as O
from mtldas_b200 import TrainPlan

kind = sys.argv[1] if len(sys.argv) > 1 else "MTL"
precision = sys.argv[2] if len(sys.argv) > 2 else "fp32"
n = int(sys.argv[3]) if len(sys.argv) > 3 else 6
sd = O.synth_state_dict(kind, seed=0)
x, yd, ye = O.synth_patches(n, seed=1234)
outs_r, loss_r, grads_r, stats_r, feats_r, fgr = O.loss_and_grads(sd, x, yd, ye, kind, dtype=torch.float64, return_feats=True)
plan = TrainPlan(kind, max(n, 8), precision, "cuda:0")
state = [v.cuda().contiguous() for v in sd.values()]
keys = list(sd.keys())
xc = x.cuda()
outs = plan.forward(state, xc)
torch.cuda.synchronize()
rel = lambda a, b: float((a.double().cpu() - b).norm() / (b.norm() + 1e-30))
for i, (o, r) in enumerate(zip(outs, outs_r)):
    print(f"out{i} rel-L2 {rel(o, r):.3e} maxabs {float((o.double().cpu()-r).abs().max()):.3e}")
for k in ("conv1.1.running_mean", "conv1.1.running_var", "resblock3.shortcut.1.running_var", "att_mask_generator4.0.4.running_mean"):
    if k in stats_r:
        print("stat", k, f"{rel(state[keys.index(k)], stats_r[k]):.3e}")
print("nbt", int(state[keys.index('conv1.1.num_batches_tracked')]), int(stats_r['conv1.1.num_batches_tracked']))
for name in ["stem"] + [f"sb{i}" for i in range(8)]:
    print("feat", name, f"{rel(plan.read_tensor(name, n), feats_r[name]):.3e}")
ntask = len(outs)
a4 = plan.read_tensor("a4", n).double().cpu()
for t in range(ntask):
    print("feat a4_%d" % t, f"{rel(a4[:, 128*t:128*(t+1)], feats_r['a4_%d' % t]):.3e}")
# backward with the reference loss gradient: d loss / d logp = -onehot / n
labels = O.task_labels(kind, yd, ye)
douts = []
for o, y in zip(outs, labels):
    d = torch.zeros_like(o)
    d[torch.arange(n), y.cuda()] = -1.0 / n
    douts.append(d)
flat = plan.backward(douts)
torch.cuda.synchronize()
for name in ["sb7", "sb6", "sb5", "sb4", "sb3", "sb2", "sb1", "sb0", "stem"]:
    try:
        got = plan.read_tensor('d:' + name, n).double().cpu()
        ref = fgr[name]
        err = (got - ref).abs()
        big = float((err > 1e-3 * ref.abs().max()).double().mean())
        print("dfeat", name, f"{rel(got, ref):.3e} max|err|/max|ref| {float(err.max()/ref.abs().max()):.3e} frac(err>1e-3 max) {big:.2e} "
              f"mean(got)/mean(ref) {float(got.sum()/ref.sum()):.6f}")
    except Exception as e:
        print("dfeat", name, "ERR", e)
off = 0
num = den = 0.0
worst = []
for k in O.param_names(kind):
    g = grads_r[k]
    m = flat[off:off + g.numel()].view(g.shape).double().cpu()
    off += g.numel()
    e = float((m - g).norm()); r = float(g.norm())
    num += e * e; den += r * r
    worst.append((e / (r + 1e-12), k, r))
print("global grad rel-L2", (num / den) ** 0.5, "launches", plan.last_launch_count)
_, _, g32, _ = O.loss_and_grads(sd, x, yd, ye, kind, dtype=torch.float32)
n2 = sum(float((g32[k].double() - grads_r[k]).norm()) ** 2 for k in grads_r)
print("oracle fp32 vs fp64 global grad rel-L2", (n2 / den) ** 0.5)
worst.sort(reverse=True)
for w in worst[:12]:
    print("  worst %.3e %s (norm %.3e)" % w)
