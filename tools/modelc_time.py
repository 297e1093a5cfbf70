import sys, os
sys.path[:0] = [".", "mtl-das.pytorch_b200"]
import torch
from mtldas_b200 import synth
from mtldas_b200.nets_c import Multi_Classifier
m = Multi_Classifier(init_weights=False); m.load_state_dict(synth.synth_state_dict_c(0), strict=True); m = m.cuda().eval()
out = []
for n in (512, 2048):
    x = synth.synth_patches(64, seed=1)[0].cuda().repeat((n + 63) // 64, 1, 1, 1)[:n].contiguous()
    for _ in range(3): m(x)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(10): m(x)
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 10
    out.append(f"batch {n}: {ms:.3f} ms {n / ms * 1e3:.0f}/s")
print(" | ".join(out))
