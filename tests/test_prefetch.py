"""Host logic of the input-side prefetcher (mtldas_b200/prefetch.py): order, values, dtypes,
ragged last batch, pass-through elements, error propagation, early exit, re-iteration -- on the CPU (plain staging); the
pinned / copy-stream path is covered by the gpu test below, fed by the reference's own file format (.mat, key 'data')."""
import os
import sys
import time

import numpy as np
import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "mtl-das.pytorch_b200")]
from mtldas_b200.prefetch import DevicePrefetcher, install_dataloader_prefetch  # noqa: E402


def _batches(n, bs, seed=0):
    g = torch.Generator().manual_seed(seed)
    x = torch.randn(n, 1, 10, 25, generator=g)
    yd, ye = torch.randint(0, 16, (n,), generator=g), torch.randint(0, 2, (n,), generator=g)
    return [(x[i:i + bs], yd[i:i + bs], ye[i:i + bs]) for i in range(0, n, bs)]


@pytest.mark.parametrize("depth", [1, 2, 4])
def test_prefetcher_keeps_order_values_and_ragged_last_batch(depth):
    ref = _batches(70, 16)                      # 16, 16, 16, 16, 6
    pf = DevicePrefetcher(ref, device="cpu", depth=depth)
    assert len(pf) == 5
    got = list(pf)
    assert len(got) == 5 and pf.batches_staged == 5
    for (x, yd, ye), (rx, ryd, rye) in zip(got, ref):
        assert isinstance(x, torch.Tensor) and x.dtype == torch.float32 and yd.dtype == torch.int64
        assert torch.equal(x, rx) and torch.equal(yd, ryd) and torch.equal(ye, rye)
    # staging slots are reused: earlier batches must not have been overwritten by later ones
    assert torch.equal(got[0][0], ref[0][0]) and torch.equal(got[1][0], ref[1][0])
    assert [tuple(b[0].shape) for b in list(pf)] == [tuple(b[0].shape) for b in ref]      # a second epoch


def test_prefetcher_passes_other_elements_through_and_accepts_numpy_and_single_tensors():
    data = [(np.full((2, 3), i, dtype=np.float32), "name%d" % i, i) for i in range(5)]
    got = list(DevicePrefetcher(data, device="cpu", depth=2))
    for i, (a, name, k) in enumerate(got):
        assert isinstance(a, torch.Tensor) and float(a.mean()) == i and name == "name%d" % i and k == i
    singles = [torch.full((4,), float(i)) for i in range(3)]
    got = list(DevicePrefetcher(singles, device="cpu", depth=2))
    assert all(isinstance(t, torch.Tensor) and float(t[0]) == i for i, t in enumerate(got))
    lists = [[torch.zeros(2), torch.ones(2)]]
    assert isinstance(list(DevicePrefetcher(lists, device="cpu"))[0], list)


def test_prefetcher_runs_ahead_propagates_errors_and_stops_early():
    class Slow:
        def __init__(self):
            self.produced = 0

        def __iter__(self):
            for i in range(6):
                self.produced += 1
                yield (torch.full((2,), float(i)),)

        def __len__(self):
            return 6

    src = Slow()
    it = iter(DevicePrefetcher(src, device="cpu", depth=2))
    first = next(it)
    time.sleep(0.3)                                   # the worker fills the queue while the consumer "computes"
    assert float(first[0][0]) == 0.0 and 3 <= src.produced <= 4      # one consumed + depth queued (+ one in hand)
    it.close()                                        # early exit: the worker must stop instead of draining the source
    time.sleep(0.3)
    assert src.produced <= 5

    def bad():
        yield (torch.zeros(1),)
        raise RuntimeError("broken file")

    it = iter(DevicePrefetcher(bad(), device="cpu", depth=2, length=2))
    next(it)
    with pytest.raises(RuntimeError, match="broken file"):
        next(it)
    with pytest.raises(ValueError):
        DevicePrefetcher([], depth=0)


def test_install_dataloader_prefetch_is_transparent_without_cuda_and_restores():
    from torch.utils.data import DataLoader, TensorDataset
    ds = TensorDataset(torch.arange(10.0).view(10, 1), torch.arange(10))
    orig = DataLoader.__iter__
    restore = install_dataloader_prefetch()
    try:
        assert DataLoader.__iter__ is not orig
        out = [b for b in DataLoader(ds, batch_size=4)]
        assert [len(b[0]) for b in out] == [4, 4, 2] and float(out[2][0][1]) == 9.0
    finally:
        restore()
    assert DataLoader.__iter__ is orig


@pytest.mark.gpu
def test_prefetcher_uploads_mat_files_through_pinned_memory(tmp_path):
    """The reference's on-disk format (dataset_preparation.py:242-273: MATLAB v5 .mat, key 'data', [100, 250] -> [1, 100,
    250] fp32, tree <root>/<k>m/*.mat, distance label int(dirname[:-1])) read by a torch DataLoader and uploaded ahead of
    the loop: device tensors, same values / labels / order as the synchronous path, `.cuda()` is a no-op on them."""
    import scipy.io as sio
    from torch.utils.data import DataLoader, Dataset
    rng = np.random.default_rng(0)
    files = []
    for k in (0, 3, 12):
        d = tmp_path / ("%dm" % k)
        d.mkdir()
        for j in range(5):
            path = str(d / ("s%d.mat" % j))
            sio.savemat(path, {"data": rng.standard_normal((100, 250))})
            files.append((path, k))

    class MatFiles(Dataset):                          # what DatasetDisk does (dataset_preparation.py:276-300)
        def __len__(self):
            return len(files)

        def __getitem__(self, i):
            mat = sio.loadmat(files[i][0])["data"][np.newaxis, :].astype(np.float32)
            return mat, files[i][1], 0

    loader = DataLoader(MatFiles(), batch_size=4, shuffle=False)
    ref = [b for b in loader]
    pf = DevicePrefetcher(loader, depth=2)
    n = 0
    for (x, yd, ye), (rx, ryd, rye) in zip(pf, ref):
        assert x.is_cuda and yd.is_cuda and x.dtype == torch.float32 and tuple(x.shape[1:]) == (1, 100, 250)
        assert x.cuda().data_ptr() == x.data_ptr()
        assert torch.equal(x.cpu(), rx) and torch.equal(yd.cpu(), ryd) and torch.equal(ye.cpu(), rye)
        n += 1
    assert n == len(ref) == 4
    restore = install_dataloader_prefetch()
    try:
        got = [b for b in DataLoader(MatFiles(), batch_size=4)]
        assert all(b[0].is_cuda for b in got) and torch.equal(got[3][0].cpu(), ref[3][0])
    finally:
        restore()


@pytest.mark.gpu
def test_prefetched_batches_train_like_synchronous_uploads():
    """The reference's loop body (utils.py:346-374: upload, forward, summed NLL, backward, Adam) over prefetched batches
    gives the same losses as over synchronously uploaded ones: the copy-stream / event hand-over is race free."""
    from model.modelA_MTL import MTL_Net
    from oracle import mtl_oracle as O
    x, yd, ye = O.synth_patches(48, seed=5)
    batches = [(x[i:i + 16], yd[i:i + 16], ye[i:i + 16]) for i in range(0, 48, 16)]
    sd = O.synth_state_dict("MTL", seed=0)
    losses = []
    for mode in ("sync", "prefetch"):
        model = MTL_Net().cuda()
        model.load_state_dict(sd, strict=True)
        model.train(True)
        opt = torch.optim.Adam(model.parameters(), lr=1e-3, weight_decay=1e-5)
        crit = torch.nn.NLLLoss()
        run = []
        for epoch in range(2):
            src = batches if mode == "sync" else DevicePrefetcher(batches, depth=2)
            for bx, d, e in src:
                bx, d, e = bx.cuda(), d.cuda(), e.cuda()
                o1, o2 = model(bx)
                loss = crit(o1, d.long()) + crit(o2, e.long())
                opt.zero_grad()
                loss.backward()
                opt.step()
                run.append(float(loss.detach()))
        losses.append(run)
    print("losses sync / prefetch:", losses)
    for a, b in zip(*losses):
        assert abs(a - b) <= 2e-2 * max(1.0, abs(a)), losses      # bf16 training steps: run-to-run atomics only
    assert losses[0][-1] < losses[0][0]
