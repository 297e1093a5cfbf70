"""One eager (no CUDA graph) MTL forward at the bench batch, for `ncu --set full` captures:
   MTLDAS_NO_GRAPH=1 ncu --set full --clock-control none --import-source on -k regex:conv_umma \
       --launch-skip 30 -c 30 -o gpurun_out/prof python tools/one_forward.py 4096
The first forward (30 conv launches) is warm-up and skipped; the second is captured."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "mtl-das.pytorch_b200")]
os.environ.setdefault("MTLDAS_NO_GRAPH", "1")

import torch  # noqa: E402
from mtldas_b200 import EvalPlan  # noqa: E402
from mtldas_b200.synth import synth_state_dict, synth_patches  # noqa: E402


def main():
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
    dev = torch.device("cuda:0")
    plan = EvalPlan("MTL", n, "bf16", dev)
    sd = synth_state_dict("MTL", seed=0)
    plan.pack({k: v.to(dev) for k, v in sd.items()})
    x = synth_patches(64, seed=1)[0].to(dev).repeat((n + 63) // 64, 1, 1, 1)[:n].contiguous()
    outs = [torch.empty(n, c, device=dev) for c in plan.classes]
    for _ in range(2):
        plan.forward(x, outs)
    torch.cuda.synchronize()
    print("ok", float(outs[0].sum()), plan.last_launch_count)


if __name__ == "__main__":
    main()
