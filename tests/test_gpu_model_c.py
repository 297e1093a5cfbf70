"""Model C (Multi_Classifier, Inception-V3 with a 1-channel stem) on the GPU, through the C ABI, against the oracle
(oracle/modelc_oracle.py, pinned to the unmodified reference by tests/golden/multi_classifier.npz)."""
import os

import pytest
import torch
import torch.nn.functional as F

from oracle import modelc_oracle as C
from oracle import mtl_oracle as O

pytestmark = pytest.mark.gpu

# (cin, cout, kh, kw, sh, sw, ph, pw, H, W): every convolution family of the model (modelC:63-83, inception.py)
CONVC_CASES = [
    (32, 32, 3, 3, 1, 1, 0, 0, 49, 124),      # Conv2d_2a_3x3 (valid)
    (32, 64, 3, 3, 1, 1, 1, 1, 47, 122),      # Conv2d_2b_3x3
    (64, 80, 1, 1, 1, 1, 0, 0, 23, 60),       # Conv2d_3b_1x1 (N = 80)
    (80, 192, 3, 3, 1, 1, 0, 0, 23, 60),      # Conv2d_4a_3x3 (five K steps: ragged last ring unit)
    (48, 64, 5, 5, 1, 1, 2, 2, 10, 28),       # InceptionA 5x5
    (288, 384, 3, 3, 2, 2, 0, 0, 10, 28),     # InceptionB 3x3 stride 2 (two N tiles of 192)
    (128, 128, 1, 7, 1, 1, 0, 3, 4, 13),      # InceptionC 1x7
    (160, 192, 7, 1, 1, 1, 3, 0, 4, 13),      # InceptionC 7x1
    (192, 320, 3, 3, 2, 2, 0, 0, 4, 13),      # InceptionD (N tiles of 160)
    (384, 384, 1, 3, 1, 1, 0, 1, 1, 6),       # InceptionE 1x3
    (448, 384, 3, 3, 1, 1, 1, 1, 1, 6),       # InceptionE 3x3 (28 K steps per tap)
    (2048, 320, 1, 1, 1, 1, 0, 0, 1, 6),      # InceptionE 1x1 over 2048 channels
    (384, 384, 3, 1, 1, 1, 1, 0, 1, 6),       # InceptionE 3x1 on one-row images (only the centre kernel row meets data)
    (192, 192, 7, 1, 1, 1, 3, 0, 4, 13, 67),  # 7x1, 67 images: every CTA of the persistent kernel walks several tiles
    (32, 64, 3, 3, 1, 1, 1, 1, 47, 122, 48),  # Conv2d_2b at 48 images: ~20 tiles per CTA (TMEM sets, rings wrap many times)
    (16, 48, 3, 3, 1, 1, 1, 1, 9, 11, 3),     # one K step only, N = 48 (16-column epilogue tail), tiny images
    (32, 32, 3, 3, 1, 1, 0, 0, 49, 124, 40),  # Conv2d_2a at 40 images: four m tiles per work item (N = 32)
    (96, 96, 3, 3, 1, 1, 1, 1, 10, 28, 300),  # InceptionA 3x3dbl_3 at 300 images: two m tiles per work item (N = 96)
]


def _planar(x):          # [n, C, H, W] fp32 -> planar-8 bf16 [C/8, n*H*W, 8]
    n, c, h, w = x.shape
    return x.permute(1, 0, 2, 3).reshape(c // 8, 8, n * h * w).permute(0, 2, 1).contiguous().to(torch.bfloat16)


@pytest.mark.parametrize("case", CONVC_CASES, ids=lambda c: "c%d-%d_k%dx%d_s%d_p%d%d_%dx%d" % (c[0], c[1], c[2], c[3], c[4], c[6], c[7], c[8], c[9]))
def test_model_c_conv_unit_matches_torch(case):
    import ctypes
    from mtldas_b200 import lib
    from mtldas_b200._lib import check
    cin, cout, kh, kw, sh, sw, ph, pw, H, W = case[:10]
    n = case[10] if len(case) > 10 else 5
    g = torch.Generator().manual_seed(cin + cout + kh)
    x = torch.randn(n, cin, H, W, generator=g)
    w = torch.randn(cout, cin, kh, kw, generator=g) / (cin * kh * kw) ** 0.5
    gamma, beta = torch.rand(cout, generator=g) + 0.5, torch.randn(cout, generator=g) * 0.2
    rm, rv = torch.randn(cout, generator=g) * 0.1, torch.rand(cout, generator=g) + 0.5
    xr = x.bfloat16().float()
    sc = gamma / torch.sqrt(rv + 1e-3)
    wf = (w * sc.view(-1, 1, 1, 1)).bfloat16().float()                    # the kernel's operands: folded, bf16
    ref = torch.relu(F.conv2d(xr.double(), wf.double(), (beta - rm * sc).double(), stride=(sh, sw), padding=(ph, pw)))
    Ho, Wo = ref.shape[2:]
    dev = torch.device("cuda:0")
    xp = _planar(xr).to(dev)
    packed = torch.empty(kh * kw * cin * cout * 2, dtype=torch.uint8, device=dev)
    bias = torch.empty(cout, device=dev)
    ctot, c0 = cout + 16, 8                                                # written at a channel offset of a wider tensor
    out = torch.zeros((ctot // 8, n * Ho * Wo, 8), dtype=torch.bfloat16, device=dev)
    p = lambda t: ctypes.c_void_p(t.data_ptr())  # noqa: E731
    st = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
    keep = [t.to(dev).contiguous() for t in (w, gamma, beta, rm, rv)]
    check(lib.mtldas_c_pack_conv(p(keep[0]), p(keep[1]), p(keep[2]), p(keep[3]), p(keep[4]), 1e-3, cout, cin, kh, kw, p(packed),
                                 p(bias), st), "pack")
    check(lib.mtldas_c_conv(p(xp), n, cin, H, W, p(packed), p(bias), p(out), ctot, c0, cout, kh, kw, sh, sw, ph, pw, 1, st), "conv")
    got = torch.empty((n, cout, Ho, Wo), device=dev)
    check(lib.mtldas_c_read(p(out), n, ctot, c0, cout, Ho * Wo, p(got), st), "read")
    torch.cuda.synchronize()
    err = float((got.cpu().double() - ref).abs().max() / ref.abs().max())
    assert err <= 6e-3, (case, err)                                        # one bf16 rounding of the stored result
    assert float(out[0].abs().max()) == 0.0 and float(out[-1].abs().max()) == 0.0      # neighbours of the slice untouched


def _fuzz_cases(count=int(os.environ.get("MTLDAS_FUZZ_CASES", "48")), seed=int(os.environ.get("MTLDAS_FUZZ_SEED", "20261107"))):
    """Random geometries the Inception program does not contain: every kernel shape with every padding from 0 to k // 2
    (also asymmetric), images down to one row / one column, strides 1 and 2, channel counts that split into one to three
    output-channel tiles, batch sizes that leave ragged last tiles and give one to several m tiles per work item."""
    import random
    rnd = random.Random(seed)
    cases = []
    while len(cases) < count:
        kh, kw = rnd.choice([(1, 1), (3, 3), (1, 3), (3, 1), (5, 5), (1, 7), (7, 1), (3, 5), (2, 2)])
        s = rnd.choice([1, 1, 1, 2])
        ph, pw = rnd.randint(0, kh // 2), rnd.randint(0, kw // 2)
        H, W = rnd.randint(1, 14), rnd.randint(1, 30)
        if H + 2 * ph < kh or W + 2 * pw < kw:
            continue
        cin = rnd.choice([16, 32, 48, 64, 80, 96, 160])
        cout = rnd.choice([16, 32, 48, 64, 96, 128, 192, 320, 384])
        n = rnd.choice([1, 2, 7, 33, 70, 150])
        cases.append((cin, cout, kh, kw, s, s, ph, pw, H, W, n))
    return cases


@pytest.mark.parametrize("case", _fuzz_cases(), ids=lambda c: "c%d-%d_k%dx%d_s%d_p%d%d_%dx%d_n%d" % (c[0], c[1], c[2], c[3], c[4], c[6], c[7], c[8], c[9], c[10]))
def test_model_c_conv_fuzz_matches_torch(case):
    test_model_c_conv_unit_matches_torch(case)


@pytest.mark.parametrize("case", [(64, 10, 28, 1, 1, 1), (768, 4, 13, 1, 1, 1), (2048, 1, 6, 1, 1, 1), (24, 5, 1, 1, 1, 1),
                                  (64, 47, 122, 0, 2, 0), (192, 21, 58, 0, 2, 0), (288, 10, 28, 0, 2, 0)],
                         ids=lambda c: "c%d_%dx%d_%s" % (c[0], c[1], c[2], "avg" if c[3] else "max"))
def test_model_c_pool_unit_matches_torch(case):
    """F.avg_pool2d(x, 3, 1, 1) (the padding counts: divide by 9 everywhere) and the 3x3 / stride-2 max pools, on bf16
    values: the average is one bf16 rounding of the fp32 sum, the maximum is exact."""
    import ctypes
    from mtldas_b200 import lib
    from mtldas_b200._lib import check
    C, H, W, is_avg, stride, pad = case
    n = 7
    x = torch.randn(n, C, H, W, generator=torch.Generator().manual_seed(C + H)).bfloat16().float()
    ref = F.avg_pool2d(x.double(), 3, 1, 1) if is_avg else F.max_pool2d(x.double(), 3, stride)
    Ho, Wo = ref.shape[2:]
    dev = torch.device("cuda:0")
    xp = _planar(x).to(dev)
    ctot, c0 = C + 16, 8
    out = torch.zeros((ctot // 8, n * Ho * Wo, 8), dtype=torch.bfloat16, device=dev)
    p = lambda t: ctypes.c_void_p(t.data_ptr())  # noqa: E731
    st = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
    check(lib.mtldas_c_pool(p(xp), n, C, H, W, p(out), ctot, c0, stride, pad, is_avg, st), "pool")
    got = torch.empty((n, C, Ho, Wo), device=dev)
    check(lib.mtldas_c_read(p(out), n, ctot, c0, C, Ho * Wo, p(got), st), "read")
    torch.cuda.synchronize()
    got = got.cpu().double()
    if is_avg:
        assert float((got - ref).abs().max()) <= 2.0 ** -8 * float(ref.abs().max())
    else:
        assert torch.equal(got, ref)
    assert float(out[0].abs().max()) == 0.0 and float(out[-1].abs().max()) == 0.0


def test_model_c_forward_matches_oracle():
    """Multi_Classifier.forward (eval) for 9 patches: logits within 1e-2 of the fp64 oracle (relative to the largest
    logit; north_star bf16 tolerance), same argmax wherever the top-2 margin is clear, stage features within 2e-2."""
    from model.modelC_multiClassifier import Multi_Classifier
    n = 9
    sd = C.synth_state_dict_c(seed=0)
    x, _, _ = O.synth_patches(n, seed=4242)
    with torch.no_grad():
        ref, feats = C.forward_c(sd, x, torch.float64, return_feats=True)
    m = Multi_Classifier(init_weights=False)
    m.load_state_dict(sd, strict=True)
    m = m.cuda().eval()
    out = m(x.cuda())
    torch.cuda.synchronize()
    assert out.shape == (n, 32) and out.dtype == torch.float32
    report = {}
    for name in ("Conv2d_1a_3x3", "Conv2d_2b_3x3", "maxpool2", "Mixed_5b", "Mixed_5d", "Mixed_6a", "Mixed_6e", "Mixed_7a", "Mixed_7c"):
        got = m.read_tensor(name).cpu().double()
        r = feats[name]
        assert got.shape == r.shape, name
        report[name] = float((got - r).norm() / r.norm())
    err = float((out.cpu().double() - ref).abs().max() / ref.abs().max())
    print("model C: stage rel-L2", {k: round(v, 5) for k, v in report.items()}, "logits max-rel", err, "launches", m.last_launch_count)
    assert all(v <= 2e-2 for v in report.values()), report
    assert err <= 1e-2, err
    top2 = ref.topk(2, dim=1).values
    clear = (top2[:, 0] - top2[:, 1]) > 4 * float((out.cpu().double() - ref).abs().max())
    assert bool((out.cpu().argmax(1) == ref.argmax(1))[clear].all())
    # a second call reuses program and packed weights; a parameter change repacks
    out2 = m(x.cuda())
    assert torch.equal(out, out2)
    with torch.no_grad():
        m.fc.bias.add_(1.0)
    out3 = m(x.cuda())
    assert float((out3 - out - 1.0).abs().max()) < 1e-5
    m.train()
    from mtldas_b200 import MtldasError
    with pytest.raises(MtldasError):
        m(x.cuda())


def test_model_c_graph_replay_equals_eager_launches():
    """forward() replays a CUDA graph captured over four streams (Inception branches side by side); the same kernels
    enqueued one after the other on one stream give the same bits, for two batch sizes and after new input data."""
    from model.modelC_multiClassifier import Multi_Classifier
    m = Multi_Classifier(init_weights=False)
    m.load_state_dict(C.synth_state_dict_c(seed=1), strict=True)
    m = m.cuda().eval()
    for n, seed in ((6, 1), (33, 2), (6, 3)):
        x = O.synth_patches(n, seed=seed)[0].cuda()
        out = m(x)
        prog = m._last_program
        eager = torch.empty_like(out)
        m._launch(prog, m._pack(x.device), x.contiguous(), eager, None)
        torch.cuda.synchronize()
        assert torch.equal(out, eager), float((out - eager).abs().max())
        assert prog.graph is not None and m.last_launch_count == len(prog.ops) + 1


def test_model_c_forward_is_batch_size_invariant():
    """The first patches of a 300-patch batch (several m tiles per work item, many items per CTA, two CTAs per SM) get the
    logits they get in a 9-patch batch (one m tile per item): the tiling never changes the order of a K reduction, and
    eval-mode BatchNorm has no batch statistics -- bit-equal rows; the 9-patch result is the one checked against the oracle."""
    from model.modelC_multiClassifier import Multi_Classifier
    m = Multi_Classifier(init_weights=False)
    m.load_state_dict(C.synth_state_dict_c(seed=0), strict=True)
    m = m.cuda().eval()
    x = O.synth_patches(300, seed=4242)[0].cuda()
    big = m(x)
    small = m(x[:9].contiguous())
    mid = m(x[:130].contiguous())
    torch.cuda.synchronize()
    assert torch.equal(big[:9], small), float((big[:9] - small).abs().max())
    assert torch.equal(big[:130], mid), float((big[:130] - mid).abs().max())
