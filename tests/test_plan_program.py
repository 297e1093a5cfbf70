"""CPU checks of the host side: the C-ABI library loads and exports its header, the plan's layer
program (interpreted in float64 by tests/p8_emulator.py) reproduces the oracle, state_dict contract."""
import os
import re

import pytest
import torch

from oracle import mtl_oracle as O

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_library_exports_every_declared_symbol():
    import ctypes
    from mtldas_b200 import lib, library_path
    hdr = open(os.path.join(ROOT, "include", "mtldas.h")).read()
    names = re.findall(r"MTLDAS_API[^;]*?\b(mtldas_\w+)\s*\(", hdr)
    assert len(names) >= 17
    raw = ctypes.CDLL(library_path())
    for n in names:
        assert hasattr(raw, n), n
    assert lib.mtldas_version().decode().startswith("mtldas-b200")


@pytest.mark.parametrize("kind", ["MTL", "single_distance", "single_event"])
def test_state_keys_match_reference_order(kind):
    from mtldas_b200 import EvalPlan
    p = EvalPlan(kind, 8, "bf16", need_device=False)
    assert p.keys == [k for k, _, _ in O.state_dict_keys(kind)]


@pytest.mark.parametrize("kind,precision", [("MTL", "bf16"), ("single_event", "fp32"), ("single_distance", "bf16")])
def test_program_reproduces_oracle(kind, precision):
    from mtldas_b200 import EvalPlan
    from p8_emulator import Emu
    nb = 3
    p = EvalPlan(kind, 4, precision, need_device=False)
    sd = O.synth_state_dict(kind, seed=0)
    x, _, _ = O.synth_patches(nb, seed=99)
    emu = Emu(p.describe(), sd, nb)
    outs = emu.run(x)
    ref = O.forward(sd, x, kind, training=False, dtype=torch.float64)
    assert len(outs) == len(ref)
    for a, b in zip(outs, ref):
        assert a.shape == b.shape
        assert float((a - b).abs().max()) < 1e-9


def test_modules_keep_reference_contract():
    from model.modelA_MTL import MTL_Net
    from model.modelB_singleTask import Single_Task_Net
    m = MTL_Net()
    assert [(k, tuple(v.shape), str(v.dtype).split(".")[1]) for k, v in m.state_dict().items()] == \
        [(k, s, d) for k, s, d in O.state_dict_keys("MTL")]
    m.load_state_dict(O.synth_state_dict("MTL"), strict=True)
    assert [n for n, _ in m.named_parameters()] == O.param_names("MTL")
    for task in ("distance", "event"):
        b = Single_Task_Net(task=task)
        assert list(b.state_dict().keys()) == [k for k, _, _ in O.state_dict_keys("single_" + task)]
    with pytest.raises(ValueError):
        Single_Task_Net(task="bogus")
    # no silent CPU fallback
    from mtldas_b200 import MtldasError
    m.eval()
    with pytest.raises(MtldasError):
        m(torch.zeros(1, 1, 100, 250))


def test_model_c_boundary_class_and_launcher():
    """Model C keeps its constructor / state_dict contract (566 keys) and refuses to run (kernels are
    the next row); the launcher binds an (absent here or present) reference utils.py to the drop-in modules."""
    from model.modelC_multiClassifier import Multi_Classifier
    from mtldas_b200 import MtldasError
    m = Multi_Classifier(init_weights=False)
    sd = m.state_dict()
    assert len(sd) == 566 and sum(p.numel() for p in m.parameters()) == 21850560
    assert tuple(sd["Conv2d_1a_3x3.conv.weight"].shape) == (32, 1, 3, 3) and tuple(sd["fc.weight"].shape) == (32, 2048)
    assert not any(k.startswith("AuxLogits") for k in sd)
    with pytest.raises(MtldasError):
        m(torch.zeros(1, 1, 100, 250))
    from mtldas_b200 import run_reference
    run_reference.prepare()
    import sys
    from mtldas_b200 import nets
    assert sys.modules["model.modelA_MTL"].MTL_Net is nets.MTL_Net
