"""Launcher for the UNMODIFIED reference scripts (train.py / test.py) on top of the drop-in modules.

    python -m mtldas_b200.run_reference /path/to/MTL-DAS.PyTorch/train.py --model MTL --GPU_device True ...

What it does before handing control to the script (runpy, `__main__`):
  * puts the drop-in `model` package (model.modelA_MTL / modelB_singleTask / modelC_multiClassifier) into
    sys.modules so `from model.modelA_MTL import MTL_Net` (utils.py:18-20) resolves to the B200 modules -- Python puts
    the script's own directory first on sys.path, so a PYTHONPATH override alone cannot shadow the reference's package;
  * provides inert stand-ins for matplotlib / seaborn / ptflops when they are not installed (utils.py:6,8,13 import
    them at module top only to draw curves and count MACs);
  * makes sklearn's confusion_matrix return an empty matrix for empty inputs (utils.py:500 crashes on sklearn >= 1.x);
  * runs every DataLoader iteration through mtldas_b200.prefetch.DevicePrefetcher (pinned staging + upload on a copy
    stream two batches ahead; the loops' own `.cuda()` calls become no-ops).  MTLDAS_PREFETCH=0 leaves the loaders alone.
`--check` only performs the set-up, imports the reference's utils.py and reports which classes it picked up.
"""
import importlib
import os
import runpy
import sys
import types

_PKG = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _stub(name, attrs=()):
    m = types.ModuleType(name)
    m.__dict__["__getattr__"] = lambda key: (lambda *a, **k: None)
    for a in attrs:
        setattr(m, a, lambda *x, **k: None)
    sys.modules[name] = m
    return m


def prepare():
    if _PKG not in sys.path:
        sys.path.insert(0, _PKG)
    for name in ("model", "model.modelA_MTL", "model.modelB_singleTask", "model.modelC_multiClassifier"):
        sys.modules.pop(name, None)
    import model  # the drop-in package next to mtldas_b200
    for sub in ("modelA_MTL", "modelB_singleTask", "modelC_multiClassifier"):
        importlib.import_module("model." + sub)
    for name in ("matplotlib", "seaborn", "ptflops"):
        try:
            importlib.import_module(name)
        except Exception:
            _stub(name)
            if name == "matplotlib":
                sys.modules["matplotlib.pyplot"] = _stub("matplotlib.pyplot")
                sys.modules["matplotlib"].pyplot = sys.modules["matplotlib.pyplot"]
    try:
        import numpy as np
        import sklearn.metrics as skm
        _orig = skm.confusion_matrix

        def confusion_matrix(y_true, y_pred, *a, **k):
            if len(y_true) == 0:
                return np.zeros((0, 0), dtype=np.int64)
            return _orig(y_true, y_pred, *a, **k)

        skm.confusion_matrix = confusion_matrix
    except Exception:
        pass
    if os.environ.get("MTLDAS_PREFETCH", "1") != "0":
        from .prefetch import install_dataloader_prefetch
        install_dataloader_prefetch(depth=2)
    return model


def main(argv=None):
    argv = list(sys.argv[1:] if argv is None else argv)
    check = "--check" in argv
    if check:
        argv.remove("--check")
    if not argv:
        raise SystemExit(__doc__)
    script = os.path.abspath(argv[0])
    prepare()
    sys.path.insert(0, os.path.dirname(script))
    if check:
        utils = importlib.import_module("utils")
        from mtldas_b200 import nets, nets_c
        ok = utils.MTL_Net is nets.MTL_Net and utils.Single_Task_Net is nets.Single_Task_Net and \
            utils.Multi_Classifier is nets_c.Multi_Classifier
        print("reference utils.py bound to the B200 drop-in modules:", ok)
        return 0 if ok else 1
    sys.argv = [script] + argv[1:]
    runpy.run_path(script, run_name="__main__")
    return 0


if __name__ == "__main__":
    raise SystemExit(main())
