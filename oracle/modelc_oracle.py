"""CPU oracle for model C (Multi_Classifier).  TEST INFRASTRUCTURE ONLY.

Functional torch-CPU restatement, over a ``state_dict``, of the eval-mode forward of

    /root/reference/model/modelC_multiClassifier.py:102-147   (Multi_Classifier._forward)
    torchvision/models/inception.py:176-365                    (InceptionA..E._forward, BasicConv2d)

i.e. Inception-V3 with a 1-channel stem (modelC:63), BatchNorm eps 1e-3 (modelC:20), 32 classes, no auxiliary head
(modelC:36,78-80), Dropout = identity in eval mode, raw logits out (the loss is CrossEntropyLoss, utils.py:137).
Pinned to the unmodified reference by tests/golden/make_golden.py --model-c (outputs committed as
tests/golden/multi_classifier_f32.npz) and re-checked by tests/test_oracle_golden.py.  The reference has no tests or
fixtures of its own for this path.  Only tests/, smoke() and bench.py's CPU legs may import this.
"""
from __future__ import annotations

from typing import Dict

import torch
import torch.nn.functional as F

from .synth import _inception_convs, state_dict_keys_c, synth_state_dict_c  # noqa: F401

BN_EPS_C = 1e-3
_SPECS = {name: (cin, cout, k, s, p) for name, cin, cout, k, s, p in _inception_convs()}


def _bc(sd, x, name, dtype):
    """BasicConv2d.forward (modelC:22-25): conv (no bias) -> BatchNorm (running statistics) -> ReLU."""
    cin, cout, k, s, p = _SPECS[name]
    y = F.conv2d(x, sd[name + ".conv.weight"].to(dtype), None, stride=s, padding=p)
    y = F.batch_norm(y, sd[name + ".bn.running_mean"].to(dtype), sd[name + ".bn.running_var"].to(dtype),
                     sd[name + ".bn.weight"].to(dtype), sd[name + ".bn.bias"].to(dtype), False, 0.1, BN_EPS_C)
    return F.relu(y)


def _inception_a(sd, x, n, d):      # torchvision InceptionA._forward
    b1 = _bc(sd, x, n + ".branch1x1", d)
    b5 = _bc(sd, _bc(sd, x, n + ".branch5x5_1", d), n + ".branch5x5_2", d)
    b3 = _bc(sd, _bc(sd, _bc(sd, x, n + ".branch3x3dbl_1", d), n + ".branch3x3dbl_2", d), n + ".branch3x3dbl_3", d)
    bp = _bc(sd, F.avg_pool2d(x, kernel_size=3, stride=1, padding=1), n + ".branch_pool", d)
    return torch.cat([b1, b5, b3, bp], 1)


def _inception_b(sd, x, n, d):      # InceptionB._forward
    b3 = _bc(sd, x, n + ".branch3x3", d)
    bd = _bc(sd, _bc(sd, _bc(sd, x, n + ".branch3x3dbl_1", d), n + ".branch3x3dbl_2", d), n + ".branch3x3dbl_3", d)
    return torch.cat([b3, bd, F.max_pool2d(x, kernel_size=3, stride=2)], 1)


def _inception_c(sd, x, n, d):      # InceptionC._forward
    b1 = _bc(sd, x, n + ".branch1x1", d)
    b7 = x
    for s in ("_1", "_2", "_3"):
        b7 = _bc(sd, b7, n + ".branch7x7" + s, d)
    bd = x
    for s in ("_1", "_2", "_3", "_4", "_5"):
        bd = _bc(sd, bd, n + ".branch7x7dbl" + s, d)
    bp = _bc(sd, F.avg_pool2d(x, kernel_size=3, stride=1, padding=1), n + ".branch_pool", d)
    return torch.cat([b1, b7, bd, bp], 1)


def _inception_d(sd, x, n, d):      # InceptionD._forward
    b3 = _bc(sd, _bc(sd, x, n + ".branch3x3_1", d), n + ".branch3x3_2", d)
    b7 = x
    for s in ("_1", "_2", "_3", "_4"):
        b7 = _bc(sd, b7, n + ".branch7x7x3" + s, d)
    return torch.cat([b3, b7, F.max_pool2d(x, kernel_size=3, stride=2)], 1)


def _inception_e(sd, x, n, d):      # InceptionE._forward
    b1 = _bc(sd, x, n + ".branch1x1", d)
    t = _bc(sd, x, n + ".branch3x3_1", d)
    b3 = torch.cat([_bc(sd, t, n + ".branch3x3_2a", d), _bc(sd, t, n + ".branch3x3_2b", d)], 1)
    t = _bc(sd, _bc(sd, x, n + ".branch3x3dbl_1", d), n + ".branch3x3dbl_2", d)
    bd = torch.cat([_bc(sd, t, n + ".branch3x3dbl_3a", d), _bc(sd, t, n + ".branch3x3dbl_3b", d)], 1)
    bp = _bc(sd, F.avg_pool2d(x, kernel_size=3, stride=1, padding=1), n + ".branch_pool", d)
    return torch.cat([b1, b3, bd, bp], 1)


def forward_c(sd: Dict[str, torch.Tensor], x: torch.Tensor, dtype=torch.float32, return_feats: bool = False):
    """Multi_Classifier.forward in eval mode (modelC:102-147,164-172): logits [N, 32]."""
    f = {}
    x = x.to(dtype)
    x = _bc(sd, x, "Conv2d_1a_3x3", dtype)
    f["Conv2d_1a_3x3"] = x
    x = _bc(sd, x, "Conv2d_2a_3x3", dtype)
    x = _bc(sd, x, "Conv2d_2b_3x3", dtype)
    f["Conv2d_2b_3x3"] = x
    x = F.max_pool2d(x, kernel_size=3, stride=2)                    # self.maxpool1 (modelC:66)
    x = _bc(sd, x, "Conv2d_3b_1x1", dtype)
    x = _bc(sd, x, "Conv2d_4a_3x3", dtype)
    x = F.max_pool2d(x, kernel_size=3, stride=2)                    # self.maxpool2 (modelC:69)
    f["maxpool2"] = x
    for n in ("Mixed_5b", "Mixed_5c", "Mixed_5d"):
        x = _inception_a(sd, x, n, dtype)
        f[n] = x
    x = _inception_b(sd, x, "Mixed_6a", dtype)
    f["Mixed_6a"] = x
    for n in ("Mixed_6b", "Mixed_6c", "Mixed_6d", "Mixed_6e"):
        x = _inception_c(sd, x, n, dtype)
        f[n] = x
    x = _inception_d(sd, x, "Mixed_7a", dtype)
    f["Mixed_7a"] = x
    for n in ("Mixed_7b", "Mixed_7c"):
        x = _inception_e(sd, x, n, dtype)
        f[n] = x
    x = torch.flatten(F.adaptive_avg_pool2d(x, (1, 1)), 1)          # avgpool + (dropout: identity in eval) + flatten
    logits = F.linear(x, sd["fc.weight"].to(dtype), sd["fc.bias"].to(dtype))
    return (logits, f) if return_feats else logits
