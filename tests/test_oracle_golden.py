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


def _sha(t):
    import hashlib
    return hashlib.sha256(t.numpy().tobytes()).hexdigest()[:16]


def test_synth_is_deterministic_and_pinned():
    """The generator is numpy-PCG64 only: the same bytes on every machine.  The stored digests pin it (a change of the
    generator invalidates the committed golden files), and the product's copy must emit the oracle's bytes."""
    import hashlib
    a = O.synth_state_dict("MTL", 0)
    x1, yd1, ye1 = O.synth_patches(3, 7)
    assert (_sha(x1), _sha(yd1), _sha(ye1)) == ("3c06952684b87093", "c6849cd02d1559f4", "85526ba79728a470")
    m = hashlib.sha256()
    for k, v in a.items():
        m.update(k.encode())
        m.update(v.numpy().tobytes())
    assert m.hexdigest()[:16] == "fc56279ec2cab93e"
    from mtldas_b200 import synth as P
    for kind in ("MTL", "single_distance", "single_event"):
        pa, oa = P.synth_state_dict(kind, 0), O.synth_state_dict(kind, 0)
        assert list(pa) == list(oa) and all(torch.equal(pa[k], oa[k]) for k in oa)
    for got, want in zip(P.synth_patches(3, 7), (x1, yd1, ye1)):
        assert torch.equal(got, want)


def test_oracle_process_loads_no_product_code():
    """bench.py --impl reference and every oracle-only process: importing the oracle must not import the product
    package, and importing the product package must not map libmtldas_b200.so (it is loaded on first use)."""
    import subprocess
    import sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    code = ("import sys; sys.path[:0] = [%r, %r]\n"
            "from oracle import mtl_oracle as O\n"
            "assert not any(m.startswith('mtldas_b200') for m in sys.modules), 'oracle imported the product'\n"
            "import mtldas_b200, mtldas_b200.synth\n"
            "assert not mtldas_b200._lib.loaded()\n"
            "assert 'libmtldas_b200' not in open('/proc/self/maps').read()\n"
            "print('clean')\n") % (root, os.path.join(root, "mtl-das.pytorch_b200"))
    r = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0 and "clean" in r.stdout, r.stderr[-2000:]


def test_model_c_oracle_matches_reference():
    """oracle/modelc_oracle.py::forward_c vs the unmodified Multi_Classifier run in the build container
    (tests/golden/multi_classifier.npz): logits and per-stage feature statistics, fp32 and fp64; 566-key contract."""
    from oracle import modelc_oracle as C
    g = np.load(os.path.join(GOLD, "multi_classifier.npz"))
    keys = C.state_dict_keys_c()
    assert len(keys) == 566 and sum(int(np.prod(s)) for k, s, d in keys if d == "float32" and "running" not in k) == 21850560
    sd = C.synth_state_dict_c(seed=0)
    x, _, _ = O.synth_patches(3, seed=4242)
    for dtype, tag, tol in ((torch.float32, "f32", 2e-4), (torch.float64, "f64", 1e-10)):
        with torch.no_grad():
            y, feats = C.forward_c(sd, x, dtype, return_feats=True)
        ref = g["logits_" + tag]
        assert y.shape == (3, 32)
        assert float(np.abs(y.numpy() - ref).max()) <= tol * max(1.0, float(np.abs(ref).max())), tag
        for key in g.files:
            if key.startswith(f"feat_{tag}:"):
                t = feats[key.split(":", 1)[1]]
                want = g[key]
                assert list(t.shape) == [int(v) for v in want[2:]], key
                assert abs(float(t.double().abs().mean()) - want[0]) <= 10 * tol * max(1.0, abs(want[0])), key
    from mtldas_b200 import synth as P
    pa = P.synth_state_dict_c(0)
    assert list(pa) == list(sd) and all(torch.equal(pa[k], sd[k]) for k in ("Conv2d_1a_3x3.conv.weight", "Mixed_6c.branch7x7dbl_3.bn.running_var", "fc.weight"))
