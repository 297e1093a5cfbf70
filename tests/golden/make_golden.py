"""Generate tests/golden/*.npz from the UNMODIFIED reference modules.

Run in the build container only (needs /root/reference):

    python tests/golden/make_golden.py

It imports /root/reference/model/modelA_MTL.py::MTL_Net and
modelB_singleTask.py::Single_Task_Net, loads the oracle's deterministic synthetic
state_dict with strict=True (which also pins the state_dict key/shape contract),
runs eval forward, train forward, the reference loss (NLLLoss per task, summed,
utils.py:139,361,367) and backward, and stores the results.  The oracle
(oracle/mtl_oracle.py) is then checked against these files by
tests/test_oracle_golden.py on any machine; the GPU path is checked against the
oracle.  Nothing at test time reads /root/reference.
"""
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, "/root/reference")

from oracle import mtl_oracle as O  # noqa: E402
from model.modelA_MTL import MTL_Net  # noqa: E402  (reference)
from model.modelB_singleTask import Single_Task_Net  # noqa: E402  (reference)

N = 6


def build(kind):
    if kind == "MTL":
        return MTL_Net()
    return Single_Task_Net(task=kind.split("_")[1])


def run(kind, dtype):
    torch.manual_seed(0)
    torch.set_num_threads(8)
    sd = O.synth_state_dict(kind, seed=0)
    x, yd, ye = O.synth_patches(N, seed=1234)
    m = build(kind)
    assert [k for k in m.state_dict().keys()] == [k for k, _, _ in O.state_dict_keys(kind)], "key order"
    m.load_state_dict(sd, strict=True)
    if dtype == torch.float64:
        m = m.double()
    out = {}
    # eval forward (utils.py:249-272)
    m.eval()
    with torch.no_grad():
        o = m(x.to(dtype))
    o = o if isinstance(o, tuple) else (o,)
    for i, t in enumerate(o):
        out[f"eval_out{i}"] = t.numpy()
    # train forward + loss + backward (utils.py:344-373)
    m.train()
    m.zero_grad()
    o = m(x.to(dtype))
    o = o if isinstance(o, tuple) else (o,)
    crit = torch.nn.NLLLoss()
    labels = O.task_labels(kind, yd, ye)
    losses = [crit(oi, li.long()) for oi, li in zip(o, labels)]
    loss = sum(losses)
    loss.backward()
    for i, t in enumerate(o):
        out[f"train_out{i}"] = t.detach().numpy()
        out[f"loss{i}"] = losses[i].detach().numpy()
    names, norms = [], []
    for k, p in m.named_parameters():
        names.append(k)
        norms.append(float(p.grad.double().norm()))
        # keep full gradients only for small tensors and a few anchor conv weights
        if p.numel() <= 128 or k in ("conv1.0.weight", "resblock1.left.0.weight", "resblock3.shortcut.0.weight",
                                     "att_mask_generato2.0.0.weight", "output_layer1.0.0.weight"):
            out["grad:" + k] = p.grad.numpy()
    out["grad_names"] = np.array(names)
    out["grad_norms"] = np.array(norms, dtype=np.float64)
    new_sd = m.state_dict()
    for k in ("conv1.1", "resblock3.shortcut.1", "att_mask_generator4.0.4", "output_layer3.0.1"):
        for leaf in ("running_mean", "running_var", "num_batches_tracked"):
            out[f"stat:{k}.{leaf}"] = new_sd[f"{k}.{leaf}"].numpy()
    return out


def main():
    for kind in ("MTL", "single_distance", "single_event"):
        for dtype, tag in ((torch.float32, "f32"), (torch.float64, "f64")):
            if kind != "MTL" and tag == "f64":
                continue
            res = run(kind, dtype)
            path = os.path.join(HERE, f"{kind}_{tag}.npz")
            np.savez_compressed(path, **res)
            print(path, os.path.getsize(path) // 1024, "KiB", "loss", [float(res[k]) for k in res if k.startswith("loss")])


if __name__ == "__main__" and "--trained" not in sys.argv and "--model-c" not in sys.argv:
    main()


def make_trained_state(steps=40, batch=32):
    """the weight state "after 40 reference Adam steps" -- torch.manual_seed(0), the
    reference constructor's default init, Adam(lr 1e-3 / 1.5, weight_decay 1e-5) as utils.py:133-134,247, batches of 32
    synthetic patches (seed 1000 + step), NLLLoss + NLLLoss (utils.py:139,361,367).  Stored as fp32 arrays in
    state_dict order (MTL_trained40.npz) with the per-step loss curve; tests/test_oracle_golden.py re-runs the same 40
    steps on the oracle restatement and must land on the same curve."""
    torch.manual_seed(0)
    torch.set_num_threads(8)
    m = MTL_Net()
    opt = torch.optim.Adam(m.parameters(), lr=1e-3 / 1.5, weight_decay=1e-5)
    crit = torch.nn.NLLLoss()
    init = {k: v.clone().numpy() for k, v in m.state_dict().items()}
    m.train()
    curve = []
    for step in range(steps):
        x, yd, ye = O.synth_patches(batch, seed=1000 + step)
        out1, out2 = m(x)
        loss = crit(out1, yd) + crit(out2, ye)
        opt.zero_grad()
        loss.backward()
        opt.step()
        curve.append(float(loss.detach()))
    out = {"key:" + k: v.numpy() for k, v in m.state_dict().items()}
    out["loss_curve"] = np.array(curve)
    # the initial state is kept in half precision only as a cross-check of the constructor's seeded init
    out["init_checksum"] = np.array([float(np.abs(v).sum()) for v in init.values()], dtype=np.float64)
    path = os.path.join(HERE, "MTL_trained40.npz")
    np.savez_compressed(path, **out)
    print(path, os.path.getsize(path) // 1024, "KiB", "loss", curve[0], "->", curve[-1])


if __name__ == "__main__" and "--trained" in sys.argv:
    make_trained_state()


def make_model_c():
    """Reference Multi_Classifier (modelC_multiClassifier.py, torchvision Inception blocks), eval mode, on the oracle's
    synthetic model-C weights (load_state_dict strict=True pins the 566-key contract): logits in fp32 and fp64 plus
    per-stage feature checksums -> tests/golden/multi_classifier.npz."""
    from model.modelC_multiClassifier import Multi_Classifier
    from oracle import modelc_oracle as C
    torch.set_num_threads(8)
    sd = C.synth_state_dict_c(seed=0)
    x, _, _ = O.synth_patches(3, seed=4242)
    out = {}
    for dtype, tag in ((torch.float32, "f32"), (torch.float64, "f64")):
        m = Multi_Classifier(init_weights=False)
        assert list(m.state_dict().keys()) == [k for k, _, _ in C.state_dict_keys_c()]
        m.load_state_dict(sd, strict=True)
        m = m.to(dtype).eval()
        feats = {}
        hooks = [getattr(m, n).register_forward_hook(lambda mod, i, o, n=n: feats.__setitem__(n, o.detach()))
                 for n in ("Conv2d_1a_3x3", "Conv2d_2b_3x3", "Mixed_5d", "Mixed_6a", "Mixed_6e", "Mixed_7a", "Mixed_7c")]
        with torch.no_grad():
            y = m(x.to(dtype))
        for h in hooks:
            h.remove()
        out["logits_" + tag] = y.numpy()
        for n, t in feats.items():
            out[f"feat_{tag}:{n}"] = np.array([float(t.double().abs().mean()), float(t.double().sum())] + list(t.shape), dtype=np.float64)
    path = os.path.join(HERE, "multi_classifier.npz")
    np.savez_compressed(path, **out)
    print(path, os.path.getsize(path), "bytes", out["logits_f32"][0, :4])


if __name__ == "__main__" and "--model-c" in sys.argv:
    make_model_c()
