"""mtldas_conv2d_nchw_static -- the convolution entry point for callers that cannot allocate GPU memory
(a streaming acquisition loop that must avoid cudaMalloc's device-wide synchronisation, or a process whose
device memory is owned by a pool) -- must give what mtldas_conv2d_nchw gives.

Collected last on purpose (file name): if it fails, a `-x` run has already executed every other test.
It is self-contained: it passes alone in a fresh process (the first call does the per-device scratch set-up)
and after the other tests.  Two calls in a row with different shapes and epilogues check that the shared
scratch is cleared between calls (the second, smaller problem would otherwise see the first one's bytes in
its zero padding)."""
import ctypes

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def _case(seed, n, cin, cout, k, stride, H, W, epi, cin2=0):
    g = torch.Generator().manual_seed(seed)
    x = torch.randn(n, cin, H, W, generator=g)
    x2 = torch.randn(n, cin2, H, W, generator=g) if cin2 else None
    w = torch.randn(cout, cin + cin2, k, k, generator=g) / np.sqrt((cin + cin2) * k * k)
    b = torch.randn(cout, generator=g) * 0.1
    ho, wo = (H, W) if stride == 1 else ((H + 1) // 2, (W + 1) // 2)
    aux = torch.randn(n, cout, ho, wo, generator=g) if epi in (2, 3) else None
    dev = "cuda:0"
    mv = lambda t: t.to(dev).contiguous() if t is not None else None  # noqa: E731
    return dict(x=mv(x), weight=mv(w), bias=mv(b), aux=mv(aux), x2=mv(x2), stride=stride, epilogue=epi)


@pytest.mark.parametrize("precision", ["bf16", "fp32"])
def test_static_scratch_conv_matches_allocating_entry_point(precision):
    from mtldas_b200 import conv2d_nchw, conv2d_nchw_static
    # first call: 3x3 + residual on the large stage-1 geometry; second call: smaller 1x1 over two sources;
    # third: stride 2 (phase-split source) -- all back to back on the same scratch
    cases = [_case(1, 5, 16, 32, 3, 1, 33, 83, 2), _case(2, 3, 32, 16, 1, 1, 17, 42, 1, cin2=32),
             _case(3, 4, 32, 64, 3, 2, 17, 42, 1)]
    want = [conv2d_nchw(precision=precision, **c) for c in cases]
    got = []
    for c in cases:            # consecutive calls, no synchronisation in between (same stream -> ordered)
        got.append(conv2d_nchw_static(precision=precision, **c))
    torch.cuda.synchronize()
    # independent comparator (torch CPU, fp64) besides the allocating entry point
    import torch.nn.functional as F
    for c, g_ in zip(cases, got):
        rd = (lambda t: t.cpu().bfloat16().double()) if precision == "bf16" else (lambda t: t.cpu().double())
        xin = torch.cat((c["x"], c["x2"]), 1) if c["x2"] is not None else c["x"]
        k = c["weight"].shape[-1]
        y = F.conv2d(rd(xin), rd(c["weight"]), c["bias"].cpu().double(), stride=c["stride"], padding=k // 2)
        ref = torch.relu(y) if c["epilogue"] == 1 else torch.relu(y + rd(c["aux"]))
        assert float((g_.cpu().double() - ref).abs().max()) <= (6e-3 if precision == "bf16" else 1e-4) * float(ref.abs().max())
    for g_, w_ in zip(got, want):
        assert g_.shape == w_.shape
        scale = float(w_.abs().max())
        # same kernels on the same data: the tolerance of the other conv tests (bf16 store rounding / 1e-4)
        tol = 6e-3 if precision == "bf16" else 1e-4
        assert float((g_ - w_).abs().max()) <= tol * scale
    # and once more in reverse order into caller-provided outputs
    for c, w_ in zip(reversed(cases), reversed(want)):
        out = torch.empty_like(w_)
        conv2d_nchw_static(precision=precision, out=out, **c)
        torch.cuda.synchronize()
        assert float((out - w_).abs().max()) <= (6e-3 if precision == "bf16" else 1e-4) * float(w_.abs().max())


def test_static_scratch_rejects_oversize_with_clear_error():
    from mtldas_b200 import lib, MtldasError, conv2d_nchw_static
    n, c, H, W = 256, 64, 33, 83          # ~140 MB of P8 temporaries: far beyond the 32 MiB built-in scratch
    x = torch.zeros(n, c, H, W, device="cuda:0")
    w = torch.zeros(64, c, 3, 3, device="cuda:0")
    with pytest.raises(MtldasError) as ei:
        conv2d_nchw_static(x, w)
    msg = str(ei.value)
    assert "scratch" in msg and str(32 << 20) in msg
    # the refusal must not have disturbed later valid calls
    c2 = _case(7, 2, 16, 16, 3, 1, 9, 21, 1)
    from mtldas_b200 import conv2d_nchw
    a = conv2d_nchw_static(**c2)
    torch.cuda.synchronize()
    b = conv2d_nchw(**c2)
    assert float((a - b).abs().max()) <= 6e-3 * float(b.abs().max())
