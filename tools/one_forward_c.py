"""Two eager (no graph, one stream) eval forwards of Multi_Classifier, for ncu:
    MTLDAS_C_NO_GRAPH=1 ncu --set full -k regex:'convc_window|convc_umma' --launch-skip 93 -c 93 python tools/one_forward_c.py 1024"""
import os
import sys

os.environ["MTLDAS_C_NO_GRAPH"] = "1"
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "mtl-das.pytorch_b200")]
import torch  # noqa: E402
from mtldas_b200 import lib, synth  # noqa: E402
from mtldas_b200.nets_c import Multi_Classifier  # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 1024
m = Multi_Classifier(init_weights=False)
m.load_state_dict(synth.synth_state_dict_c(0), strict=True)
m = m.cuda().eval()
x = synth.synth_patches(64, seed=1)[0].cuda().repeat((n + 63) // 64, 1, 1, 1)[:n].contiguous()
for _ in range(2):
    out = m(x)
torch.cuda.synchronize()
print(lib.mtldas_version().decode(), "batch", n, "logits", float(out.abs().mean()))
