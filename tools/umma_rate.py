import ctypes, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "mtl-das.pytorch_b200")]
import torch
from mtldas_b200 import lib
torch.zeros(1, device="cuda")
out = (ctypes.c_float * 2)()
for mode in (0, 1, 2):
    for n in (16, 32, 64, 128, 256):
        for reps in (1, 4):
            rc = lib.mtldas_umma_rate(n, mode, reps, out)
            print(f"mode {mode} N={n:3d} reps={reps}: rc={rc} cycles/MMA total {out[0]:.1f} issue {out[1]:.1f}", flush=True)
