"""Streaming inference over a large DAS array, sharded by fibre-channel windows (BASELINE.json configs[4]).

The reference classifies isolated [100 channels x 250 samples] patches (utils.py:128); a deployed interrogator
produces one long array [n_channels, n_time].  The network is applied to every window of 100 channels (stride
`stride`) and every block of 250 samples.  Windows are independent, so ranks take contiguous window ranges and
nothing is exchanged: each rank needs only the channel slab of its windows (plus the window-stride halo).
"""
import ctypes

import torch

from ._lib import lib, check, MtldasError
from .ddp import shard_windows
from .engine import EvalPlan, _stream_ptr

WINDOW, TBLOCK = 100, 250


class StreamingInference:
    def __init__(self, state_dict, kind="MTL", precision="bf16", micro_batch=512, stride=10, device=None, in_place=True):
        self.in_place = in_place      # False: always gather the windows into a patch tensor first (mtldas_window_gather)
        self.plan = EvalPlan(kind, micro_batch, precision, device)
        self.device, self.stride = self.plan.device, int(stride)
        self.plan.pack({k: v.to(self.device) for k, v in state_dict.items()})

    def windows_of_rank(self, n_channels, rank=0, world=1):
        """(first channels of this rank's windows, (c_lo, c_hi) channel slab it needs)."""
        return shard_windows(n_channels, rank, world, WINDOW, self.stride)

    def run(self, slab, first_channel=0, n_windows=None):
        """slab: CUDA fp32 [channels, n_time] holding this rank's channel slab.  Returns per-task log-probs
        shaped [n_windows, n_time // 250, classes] for windows starting at first_channel + i*stride."""
        if not slab.is_cuda or slab.dtype != torch.float32 or slab.dim() != 2:
            raise MtldasError("slab must be a CUDA fp32 [channels, time] tensor")
        slab = slab.contiguous()
        n_ch, n_time = slab.shape
        if n_windows is None:
            n_windows = max(0, (n_ch - first_channel - WINDOW) // self.stride + 1)
        n_tb = n_time // TBLOCK
        n = n_windows * n_tb
        if self.in_place and n and self.plan.precision == "bf16" and n_time % 4 == 0 and slab.data_ptr() % 16 == 0:
            # the stem cuts the windows out of the array itself: no patch tensor is written or re-read
            outs = [torch.empty((n, c), dtype=torch.float32, device=self.device) for c in self.plan.classes]
            with torch.cuda.device(self.device):
                check(lib.mtldas_plan_forward_eval_windows(
                    self.plan._h, ctypes.c_void_p(slab.data_ptr()), n_ch, n_time, first_channel, self.stride, n_windows,
                    ctypes.c_void_p(outs[0].data_ptr()), ctypes.c_void_p(outs[1].data_ptr()) if len(outs) > 1 else None,
                    _stream_ptr(self.device)), "forward_eval_windows")
            return tuple(o.view(n_windows, n_tb, -1) for o in outs)
        patches = torch.empty((n, 1, WINDOW, TBLOCK), dtype=torch.float32, device=self.device)
        with torch.cuda.device(self.device):
            check(lib.mtldas_window_gather(ctypes.c_void_p(slab.data_ptr()), n_ch, n_time, first_channel, self.stride,
                                           n_windows, ctypes.c_void_p(patches.data_ptr()), _stream_ptr(self.device)),
                  "window_gather")
        outs = self.plan.forward(patches) if n else tuple(torch.empty((0, c), device=self.device) for c in self.plan.classes)
        return tuple(o.view(n_windows, n_tb, -1) for o in outs)
