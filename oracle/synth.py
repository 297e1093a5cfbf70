"""TEST INFRASTRUCTURE (oracle side).  Deterministic synthetic weights and DAS patches (numpy PCG64 uniform stream
only, so the same bytes are produced on any machine and torch version).  This is the oracle's own generator: the
product package has an equal one (mtldas_b200/synth.py) and tests/test_oracle_golden.py pins the two to each other
and to stored checksums, so that an oracle-only process (bench.py --impl reference) never imports product code."""
import math
from typing import Dict, List, Tuple

import numpy as np
import torch

PATCH_H, PATCH_W = 100, 250

_RES = [(16, 16, 1), (16, 16, 1), (16, 32, 2), (32, 32, 1), (32, 64, 2), (64, 64, 1), (64, 128, 2), (128, 128, 1)]
_ATT = ["att_mask_generator1", "att_mask_generato2", "att_mask_generator3", "att_mask_generator4"]
_ATT_CH = [(16, 8, 16), (64, 16, 32), (128, 32, 64), (256, 64, 128)]
_OUT_CH = [(16, 32), (32, 64), (64, 128)]


def _ntasks(kind: str) -> int:
    if kind == "MTL":
        return 2
    if kind in ("single_distance", "single_event"):
        return 1
    raise ValueError(kind)


def state_dict_keys(kind: str) -> List[Tuple[str, Tuple[int, ...], str]]:
    """(key, shape, dtype) of the reference state_dict, in registration order."""
    nt = _ntasks(kind)
    convs = [("conv1.0", "conv1.1", 1, 16, 7, False)]
    for i, (ci, co, s) in enumerate(_RES, start=1):
        convs.append((f"resblock{i}.left.0", f"resblock{i}.left.1", ci, co, 3, False))
        convs.append((f"resblock{i}.left.3", f"resblock{i}.left.4", co, co, 3, False))
        if s != 1 or ci != co:
            convs.append((f"resblock{i}.shortcut.0", f"resblock{i}.shortcut.1", ci, co, 1, False))
    for name, (ci, cm, co) in zip(_ATT, _ATT_CH):
        for t in range(nt):
            convs.append((f"{name}.{t}.0", f"{name}.{t}.1", ci, cm, 1, True))
            convs.append((f"{name}.{t}.3", f"{name}.{t}.4", cm, co, 3, True))
    for lvl, (ci, co) in enumerate(_OUT_CH, start=1):
        for t in range(nt):
            convs.append((f"output_layer{lvl}.{t}.0", f"output_layer{lvl}.{t}.1", ci, co, 3, False))
    keys = []
    for conv, bn, ci, co, k, bias in convs:
        keys.append((conv + ".weight", (co, ci, k, k), "float32"))
        if bias:
            keys.append((conv + ".bias", (co,), "float32"))
        keys += [(bn + ".weight", (co,), "float32"), (bn + ".bias", (co,), "float32"),
                 (bn + ".running_mean", (co,), "float32"), (bn + ".running_var", (co,), "float32"),
                 (bn + ".num_batches_tracked", (), "int64")]
    return keys


def synth_state_dict(kind: str = "MTL", seed: int = 0, trained_like: bool = True) -> Dict[str, torch.Tensor]:
    """A full state_dict with non-trivial BN affine/running statistics.

    Conv weights/biases use the bound torch's default init uses (U(+-1/sqrt(fan_in))),
    BN gamma/beta/running stats are perturbed so that eval-mode folding and the
    affine terms are really exercised (torch's defaults gamma=1,beta=0,rm=0,rv=1 hide bugs).
    """
    rng = np.random.default_rng(seed)
    sd: Dict[str, torch.Tensor] = {}

    def u(shape, lo, hi):
        return torch.from_numpy((rng.random(shape, dtype=np.float64) * (hi - lo) + lo).astype(np.float32))

    for key, shape, dtype in state_dict_keys(kind):
        leaf = key.rsplit(".", 1)[1]
        if dtype == "int64":
            sd[key] = torch.tensor(3 if trained_like else 0, dtype=torch.int64)
        elif len(shape) == 4:
            fan_in = shape[1] * shape[2] * shape[3]
            # a little hotter than kaiming-uniform(a=sqrt5) so activations do not fade with depth
            b = math.sqrt(3.0 / fan_in)
            sd[key] = u(shape, -b, b)
            last_fan_in = fan_in
        elif leaf == "bias" and key.split(".")[-2] in ("0", "3") and "generato" in key:
            b = 1.0 / math.sqrt(last_fan_in)
            sd[key] = u(shape, -b, b)
        elif leaf == "weight":        # BN gamma
            sd[key] = u(shape, 0.6, 1.4) if trained_like else torch.ones(shape)
        elif leaf == "bias":          # BN beta
            sd[key] = u(shape, -0.3, 0.3) if trained_like else torch.zeros(shape)
        elif leaf == "running_mean":
            sd[key] = u(shape, -0.2, 0.2) if trained_like else torch.zeros(shape)
        elif leaf == "running_var":
            sd[key] = u(shape, 0.5, 1.5) if trained_like else torch.ones(shape)
        else:
            raise AssertionError(key)
    return sd


def synth_patches(n: int, seed: int = 1234) -> Tuple[torch.Tensor, torch.Tensor, torch.Tensor]:
    """Synthetic DAS patches [n,1,100,250] fp32 + (distance, event) int64 labels.

    N(0,1)-like noise (sum of uniforms, numpy-only) plus a label-dependent burst pattern
    (SURVEY.md section 8d): vertical half-width 5+2*distance around row 50, 6-column bursts of
    amplitude 3 with period 60 (event 0) or 25 (event 1).
    """
    rng = np.random.default_rng(seed)
    x = (rng.random((n, 1, PATCH_H, PATCH_W, 4), dtype=np.float32).sum(-1) - 2.0) * math.sqrt(3.0)
    yd = rng.integers(0, 16, size=n, dtype=np.int64)
    ye = rng.integers(0, 2, size=n, dtype=np.int64)
    for i in range(n):
        hw = 5 + 2 * int(yd[i])
        r0, r1 = max(0, 50 - hw), min(PATCH_H, 50 + hw)
        period = 60 if ye[i] == 0 else 25
        for c0 in range(7, PATCH_W - 6, period):
            x[i, 0, r0:r1, c0:c0 + 6] += 3.0
    return torch.from_numpy(x.astype(np.float32)), torch.from_numpy(yd), torch.from_numpy(ye)


