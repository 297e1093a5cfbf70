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


def test_model_c_layer_program_shapes():
    """Host logic of the model-C layer program (nets_c._Program): 94 convolutions, the pools, concat offsets and the
    spatial bookkeeping of modelC_multiClassifier.py:102-147 for 100 x 250 patches -- bit-exact integer work."""
    from mtldas_b200.nets_c import _Program, _SPECS
    prog = _Program(2, "cpu")
    convs = [op for op in prog.ops if op[0] == "conv"]
    pools = [op for op in prog.ops if op[0] == "pool"]
    assert len(convs) == 94 and sorted(op[4] for op in convs) == sorted(_SPECS)
    assert len(pools) == 2 + 3 + 1 + 4 + 1 + 2          # maxpool1/2, avg pools of A/C/E blocks, max pools of B/D
    want = {"Conv2d_1a_3x3": (32, 49, 124), "Conv2d_2b_3x3": (64, 47, 122), "maxpool1": (64, 23, 60),
            "Conv2d_4a_3x3": (192, 21, 58), "maxpool2": (192, 10, 28), "Mixed_5b": (256, 10, 28), "Mixed_5d": (288, 10, 28),
            "Mixed_6a": (768, 4, 13), "Mixed_6e": (768, 4, 13), "Mixed_7a": (1280, 1, 6), "Mixed_7c": (2048, 1, 6)}
    for k, v in want.items():
        assert prog.shapes[k] == v, (k, prog.shapes[k])
    assert prog.last == "Mixed_7c"
    # every channel of every block output is written exactly once
    for blk in ("Mixed_5b", "Mixed_6a", "Mixed_6c", "Mixed_7a", "Mixed_7c"):
        cover = []
        for kind, src, dst, c0, arg in prog.ops:
            if dst == blk:
                width = _SPECS[arg][1] if kind == "conv" else prog.shapes[src][0]
                cover.append((c0, c0 + width))
        cover.sort()
        assert cover[0][0] == 0 and cover[-1][1] == prog.shapes[blk][0]
        assert all(a[1] == b[0] for a, b in zip(cover, cover[1:])), (blk, cover)
    # stream schedule of the captured graph: dependencies are exactly the writers of the op's source tensor, a chain of a
    # branch stays on one lane, the four branches of a block are on four different lanes, nothing is written twice
    assert len(prog.deps) == len(prog.lane) == len(prog.ops)
    for i, (kind, src, dst, c0, arg) in enumerate(prog.ops):
        writers = [j for j, op in enumerate(prog.ops) if op[2] == src]
        assert prog.deps[i] == writers and all(j < i for j in writers), (i, arg)
        if "." in src:                                   # a branch-private tensor: its consumer stays on the producer's lane
            assert len(writers) == 1 and prog.lane[i] == prog.lane[writers[0]], (i, arg)
    for blk in ("Mixed_5c", "Mixed_6d", "Mixed_7b"):
        heads = [i for i, op in enumerate(prog.ops) if op[1] == prog.ops[[o[2] for o in prog.ops].index(blk)][1]
                 and (op[2] == blk or op[2].startswith(blk + "."))]
        assert sorted(prog.lane[i] for i in heads) == [0, 1, 2, 3], (blk, [prog.lane[i] for i in heads])
    written = [(op[2], op[3]) for op in prog.ops]
    assert len(set(written)) == len(written)


def test_model_c_boundary_class_and_launcher():
    """Model C keeps its constructor / state_dict contract (566 keys) and has no CPU fallback;
    the launcher binds an (absent here or present) reference utils.py to the drop-in modules."""
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
