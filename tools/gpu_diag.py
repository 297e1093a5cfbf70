"""First-contact diagnostics for a GPU box: each probe runs in its own process (a trapped kernel kills
the CUDA context) with a timeout, prints one line, never raises.  Output goes to gpurun_out/diag.log."""
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PRE = f"import sys; sys.path[:0]=[{ROOT!r}, {os.path.join(ROOT, 'mtl-das.pytorch_b200')!r}]; import torch; "

PROBES = {
    "device": "print(torch.cuda.get_device_name(0), torch.cuda.get_device_capability(0))",
    "selftest": "from mtldas_b200 import umma_selftest as u; "
                "print({(n,s,w): u(n,s,w) for n in (16,64,128) for s in (0,1,85) for w in (0,1)})",
    "conv_fp32": "from mtldas_b200 import conv2d_nchw as c; import torch.nn.functional as F; "
                 "x=torch.randn(2,16,33,83); w=torch.randn(16,16,3,3)/12; "
                 "y=c(x.cuda(),w.cuda(),precision='fp32').cpu(); print('err',float((y-F.conv2d(x,w,padding=1)).abs().max()))",
    "conv_bf16": "from mtldas_b200 import conv2d_nchw as c; import torch.nn.functional as F; "
                 "x=torch.randn(2,16,33,83).bfloat16().float(); w=(torch.randn(16,16,3,3)/12).bfloat16().float(); "
                 "y=c(x.cuda(),w.cuda(),precision='bf16').cpu(); r=F.conv2d(x,w,padding=1); "
                 "print('err',float((y-r).abs().max()),'scale',float(r.abs().max()))",
    "conv_bf16_1x1": "from mtldas_b200 import conv2d_nchw as c; import torch.nn.functional as F; "
                 "x=torch.randn(2,16,33,83).bfloat16().float(); w=(torch.randn(16,16,1,1)/4).bfloat16().float(); "
                 "y=c(x.cuda(),w.cuda(),precision='bf16').cpu(); r=F.conv2d(x,w); "
                 "print('err',float((y-r).abs().max()),'scale',float(r.abs().max()))",
    "conv_bf16_swap": "import os; os.environ['MTLDAS_UMMA_SWAP_DESC']='1'; "
                 "from mtldas_b200 import conv2d_nchw as c; import torch.nn.functional as F; "
                 "x=torch.randn(2,16,33,83).bfloat16().float(); w=(torch.randn(16,16,1,1)/4).bfloat16().float(); "
                 "y=c(x.cuda(),w.cuda(),precision='bf16').cpu(); r=F.conv2d(x,w); "
                 "print('err',float((y-r).abs().max()),'scale',float(r.abs().max()))",
    "conv_bf16_ring": "from mtldas_b200 import conv2d_nchw as c; import torch.nn.functional as F; "
                 "x=torch.randn(3,128,5,11).bfloat16().float(); w=(torch.randn(128,128,3,3)/34).bfloat16().float(); "
                 "y=c(x.cuda(),w.cuda(),precision='bf16').cpu(); r=F.conv2d(x,w,padding=1); "
                 "print('err',float((y-r).abs().max()),'scale',float(r.abs().max()))",
}


def main():
    for name, code in PROBES.items():
        try:
            r = subprocess.run([sys.executable, "-c", PRE + code], capture_output=True, text=True, timeout=120)
            tail = (r.stdout.strip().splitlines() or [""])[-1]
            err = " | ".join(r.stderr.strip().splitlines()[-3:]) if r.returncode else ""
            print(f"[diag] {name}: rc={r.returncode} {tail} {err}", flush=True)
        except subprocess.TimeoutExpired:
            print(f"[diag] {name}: TIMEOUT", flush=True)


if __name__ == "__main__":
    main()
