"""Deterministic synthetic weights and DAS patches (numpy PCG64 uniform stream only, so the same bytes
are produced on any machine and torch version).  Used by bench.py, smoke() and the tests; the CPU oracle
re-exports these so both sides of every comparison see identical inputs."""
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


# --------------------------------------------------------------------------------------
# model C (Multi_Classifier, Inception-V3 with a 1-channel stem): layer table, state_dict contract, synthetic weights
# --------------------------------------------------------------------------------------
def _inception_convs():
    """(prefix, cin, cout, (kh, kw), (sh, sw), (ph, pw)) of every BasicConv2d in registration order
    (modelC_multiClassifier.py:63-83; torchvision/models/inception.py InceptionA..E constructors)."""
    L = []

    def add(name, cin, cout, k=1, s=1, p=0):
        k = (k, k) if isinstance(k, int) else k
        s = (s, s) if isinstance(s, int) else s
        p = (p, p) if isinstance(p, int) else p
        L.append((name, cin, cout, k, s, p))

    add("Conv2d_1a_3x3", 1, 32, 3, 2)
    add("Conv2d_2a_3x3", 32, 32, 3)
    add("Conv2d_2b_3x3", 32, 64, 3, 1, 1)
    add("Conv2d_3b_1x1", 64, 80, 1)
    add("Conv2d_4a_3x3", 80, 192, 3)
    for name, cin, pf in (("Mixed_5b", 192, 32), ("Mixed_5c", 256, 64), ("Mixed_5d", 288, 64)):      # InceptionA
        add(name + ".branch1x1", cin, 64)
        add(name + ".branch5x5_1", cin, 48)
        add(name + ".branch5x5_2", 48, 64, 5, 1, 2)
        add(name + ".branch3x3dbl_1", cin, 64)
        add(name + ".branch3x3dbl_2", 64, 96, 3, 1, 1)
        add(name + ".branch3x3dbl_3", 96, 96, 3, 1, 1)
        add(name + ".branch_pool", cin, pf)
    add("Mixed_6a.branch3x3", 288, 384, 3, 2)                                                          # InceptionB
    add("Mixed_6a.branch3x3dbl_1", 288, 64)
    add("Mixed_6a.branch3x3dbl_2", 64, 96, 3, 1, 1)
    add("Mixed_6a.branch3x3dbl_3", 96, 96, 3, 2)
    for name, c7 in (("Mixed_6b", 128), ("Mixed_6c", 160), ("Mixed_6d", 160), ("Mixed_6e", 192)):      # InceptionC
        add(name + ".branch1x1", 768, 192)
        add(name + ".branch7x7_1", 768, c7)
        add(name + ".branch7x7_2", c7, c7, (1, 7), 1, (0, 3))
        add(name + ".branch7x7_3", c7, 192, (7, 1), 1, (3, 0))
        add(name + ".branch7x7dbl_1", 768, c7)
        add(name + ".branch7x7dbl_2", c7, c7, (7, 1), 1, (3, 0))
        add(name + ".branch7x7dbl_3", c7, c7, (1, 7), 1, (0, 3))
        add(name + ".branch7x7dbl_4", c7, c7, (7, 1), 1, (3, 0))
        add(name + ".branch7x7dbl_5", c7, 192, (1, 7), 1, (0, 3))
        add(name + ".branch_pool", 768, 192)
    add("Mixed_7a.branch3x3_1", 768, 192)                                                              # InceptionD
    add("Mixed_7a.branch3x3_2", 192, 320, 3, 2)
    add("Mixed_7a.branch7x7x3_1", 768, 192)
    add("Mixed_7a.branch7x7x3_2", 192, 192, (1, 7), 1, (0, 3))
    add("Mixed_7a.branch7x7x3_3", 192, 192, (7, 1), 1, (3, 0))
    add("Mixed_7a.branch7x7x3_4", 192, 192, 3, 2)
    for name, cin in (("Mixed_7b", 1280), ("Mixed_7c", 2048)):                                         # InceptionE
        add(name + ".branch1x1", cin, 320)
        add(name + ".branch3x3_1", cin, 384)
        add(name + ".branch3x3_2a", 384, 384, (1, 3), 1, (0, 1))
        add(name + ".branch3x3_2b", 384, 384, (3, 1), 1, (1, 0))
        add(name + ".branch3x3dbl_1", cin, 448)
        add(name + ".branch3x3dbl_2", 448, 384, 3, 1, 1)
        add(name + ".branch3x3dbl_3a", 384, 384, (1, 3), 1, (0, 1))
        add(name + ".branch3x3dbl_3b", 384, 384, (3, 1), 1, (1, 0))
        add(name + ".branch_pool", cin, 192)
    return L


def state_dict_keys_c():
    """(key, shape, dtype) of the reference Multi_Classifier state_dict in registration order (566 entries)."""
    keys = []
    for name, cin, cout, k, s, p in _inception_convs():
        keys.append((name + ".conv.weight", (cout, cin, k[0], k[1]), "float32"))
        keys += [(name + ".bn.weight", (cout,), "float32"), (name + ".bn.bias", (cout,), "float32"),
                 (name + ".bn.running_mean", (cout,), "float32"), (name + ".bn.running_var", (cout,), "float32"),
                 (name + ".bn.num_batches_tracked", (), "int64")]
    keys += [("fc.weight", (32, 2048), "float32"), ("fc.bias", (32,), "float32")]
    return keys


def synth_state_dict_c(seed: int = 0):
    """Deterministic synthetic weights for model C (numpy PCG64 only): He-uniform conv weights so activations keep their
    scale through 94 ReLU layers, perturbed BatchNorm affine / running statistics (the eval fold must matter)."""
    rng = np.random.default_rng(seed)
    sd = {}

    def u(shape, lo, hi):
        return torch.from_numpy((rng.random(shape, dtype=np.float64) * (hi - lo) + lo).astype(np.float32))

    for key, shape, dtype in state_dict_keys_c():
        leaf = key.rsplit(".", 1)[1]
        if dtype == "int64":
            sd[key] = torch.tensor(3, dtype=torch.int64)
        elif len(shape) == 4:
            b = math.sqrt(6.0 / (shape[1] * shape[2] * shape[3]))
            sd[key] = u(shape, -b, b)
        elif key == "fc.weight":
            b = math.sqrt(3.0 / shape[1])
            sd[key] = u(shape, -b, b)
        elif key == "fc.bias":
            sd[key] = u(shape, -0.1, 0.1)
        elif leaf == "weight":
            sd[key] = u(shape, 0.7, 1.3)
        elif leaf == "bias":
            sd[key] = u(shape, -0.2, 0.2)
        elif leaf == "running_mean":
            sd[key] = u(shape, -0.2, 0.2)
        elif leaf == "running_var":
            sd[key] = u(shape, 0.6, 1.4)
        else:
            raise AssertionError(key)
    return sd
