"""The oracle restatement vs golden vectors produced by the unmodified reference
(tests/golden/make_golden.py).  CPU only."""
import os

import numpy as np
import pytest
import torch

from oracle import mtl_oracle as O

GOLD = os.path.join(os.path.dirname(__file__), "golden")
N = 6


def _load(kind, tag):
    return np.load(os.path.join(GOLD, f"{kind}_{tag}.npz"))


@pytest.mark.parametrize("kind,tag", [("MTL", "f32"), ("MTL", "f64"), ("single_distance", "f32"),
                                      ("single_event", "f32")])
def test_oracle_matches_reference(kind, tag):
    g = _load(kind, tag)
    dtype = torch.float32 if tag == "f32" else torch.float64
    tol = 2e-5 if tag == "f32" else 1e-10
    sd = O.synth_state_dict(kind, seed=0)
    x, yd, ye = O.synth_patches(N, seed=1234)
    outs = O.forward(sd, x, kind, training=False, dtype=dtype)
    for i, o in enumerate(outs):
        np.testing.assert_allclose(o.numpy(), g[f"eval_out{i}"], rtol=tol, atol=tol)
    outs, loss, grads, stats = O.loss_and_grads(sd, x, yd, ye, kind, dtype=dtype)
    for i, o in enumerate(outs):
        np.testing.assert_allclose(o.numpy(), g[f"train_out{i}"], rtol=tol, atol=tol)
    total = sum(float(g[f"loss{i}"]) for i in range(len(outs)))
    assert abs(float(loss) - total) <= tol * 10
    names = [str(s) for s in g["grad_names"]]
    assert names == O.param_names(kind)
    gtol = 2e-3 if tag == "f32" else 1e-8      # fp32 grads of this net are ill-conditioned
    for k, ref_norm in zip(names, g["grad_norms"]):
        mine = float(grads[k].double().norm())
        assert abs(mine - ref_norm) <= gtol * max(ref_norm, 1e-3), (k, mine, ref_norm)
    for key in g.files:
        if key.startswith("grad:"):
            ref = g[key]
            np.testing.assert_allclose(grads[key[5:]].numpy(), ref, rtol=0, atol=gtol * max(np.abs(ref).max(), 1e-4),
                                       err_msg=key)
        if key.startswith("stat:"):
            np.testing.assert_allclose(stats[key[5:]].numpy(), g[key], rtol=1e-5, atol=1e-6, err_msg=key)


def test_state_dict_contract_counts():
    # 268 keys / 1 136 224 params (A); 194 keys / 918 376 params (B)
    for kind, nkeys, nparams in (("MTL", 268, 1136224), ("single_distance", 194, 918376), ("single_event", 194, 918376)):
        keys = O.state_dict_keys(kind)
        assert len(keys) == nkeys
        pn = set(O.param_names(kind))
        assert sum(int(np.prod(s)) for k, s, d in keys if k in pn) == nparams


def test_synth_is_deterministic():
    a = O.synth_state_dict("MTL", 0)
    b = O.synth_state_dict("MTL", 0)
    assert all(torch.equal(a[k], b[k]) for k in a)
    x1, yd1, ye1 = O.synth_patches(3, 7)
    x2, yd2, ye2 = O.synth_patches(3, 7)
    assert torch.equal(x1, x2) and torch.equal(yd1, yd2) and torch.equal(ye1, ye2)
    # checksum pins the generator across machines
    assert abs(float(x1.double().sum()) - float(O.synth_patches(3, 7)[0].double().sum())) == 0
