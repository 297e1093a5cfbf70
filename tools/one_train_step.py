"""Three eager (no CUDA graph) MTL training steps (forward with batch statistics + backward) at the bench
batch, for ncu launch lists:  ncu --metrics gpu__time_duration.sum --clock-control none --csv \
    --log-file gpurun_out/train_launches.csv python tools/one_train_step.py 256"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "mtl-das.pytorch_b200")]
os.environ.setdefault("MTLDAS_NO_GRAPH", "1")

import torch  # noqa: E402
from mtldas_b200 import TrainPlan, synth as O  # noqa: E402


def main():
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 256
    steps = int(sys.argv[2]) if len(sys.argv) > 2 else 3
    sd = O.synth_state_dict("MTL", seed=0)
    x, _, _ = O.synth_patches(min(n, 64), seed=1)
    x = x.repeat((n + 63) // 64, 1, 1, 1)[:n].cuda()
    plan = TrainPlan("MTL", n, "bf16", "cuda:0")
    state = [v.cuda().contiguous() for v in sd.values()]
    d0 = torch.full((n, 16), -1.0 / n, device="cuda")
    d1 = torch.full((n, 2), -1.0 / n, device="cuda")
    flat = torch.empty(plan.n_params, device="cuda")
    for _ in range(steps):
        plan.forward(state, x)
        f = plan.last_launch_count
        plan.backward([d0, d1], flat)
        b = plan.last_launch_count
    torch.cuda.synchronize()
    print("ok launches fwd", f, "bwd", b, float(flat.abs().sum()))


if __name__ == "__main__":
    main()
