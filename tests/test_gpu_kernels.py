"""GPU parity tests proper: every call goes through the C ABI (ctypes) into the sm_100a kernels and is
compared with the CPU oracle / torch CPU functional ops on the same seeded inputs."""
import numpy as np
import pytest
import torch
import torch.nn.functional as F

from oracle import mtl_oracle as O

pytestmark = pytest.mark.gpu

# (cin, cout, k, stride, H, W, epilogue, cin2, aux_phase_split) -- the conv configurations of SURVEY App. B
CONV_CASES = [
    (16, 16, 3, 1, 33, 83, 1, 0, 0),     # resblock1/2 convs (BN+ReLU)
    (16, 16, 3, 1, 33, 83, 2, 0, 0),     # + residual add
    (16, 32, 3, 1, 33, 83, 3, 0, 1),     # mask generator 3x3: sigmoid * feature (phase-split feature)
    (16, 32, 3, 2, 33, 83, 1, 0, 0),     # resblock3.left.0 stride 2
    (16, 32, 1, 2, 33, 83, 0, 0, 0),     # resblock3.shortcut
    (32, 32, 3, 1, 17, 42, 2, 0, 0),
    (32, 16, 1, 1, 17, 42, 1, 32, 0),    # mask generator 1x1 over cat(shared, task)
    (32, 64, 3, 2, 17, 42, 1, 0, 0),
    (32, 64, 3, 1, 17, 42, 1, 0, 0),     # output_layer2
    (64, 64, 3, 1, 9, 21, 2, 0, 0),
    (64, 128, 3, 2, 9, 21, 1, 0, 0),
    (64, 128, 1, 2, 9, 21, 0, 0, 0),
    (64, 128, 3, 1, 9, 21, 1, 0, 0),     # output_layer3 (weight ring)
    (128, 128, 3, 1, 5, 11, 2, 0, 0),    # resblock8 (weight ring, 4 slots)
    (128, 64, 1, 1, 5, 11, 1, 128, 0),   # att4 1x1 256->64
    (64, 128, 3, 1, 5, 11, 3, 0, 0),     # att4 3x3 sigmoid * sb7
]


def _ref_conv(x, w, b, aux, x2, stride, epi, k):
    xin = torch.cat((x, x2), 1) if x2 is not None else x
    y = F.conv2d(xin.double(), w.double(), b.double(), stride=stride, padding=k // 2)
    if epi == 1:
        y = torch.relu(y)
    elif epi == 2:
        y = torch.relu(y + aux.double())
    elif epi == 3:
        y = torch.sigmoid(y) * aux.double()
    return y


# fused ReLU + MaxPool2d(2,2,ceil) epilogue (output_layer1/2/3): odd H and W give the clipped last row / column windows
POOL_CASES = [(16, 32, 33, 83), (32, 64, 17, 42), (64, 128, 9, 21)]


@pytest.mark.parametrize("case", POOL_CASES, ids=lambda c: "c%d-%d_%dx%d" % c)
def test_fused_relu_maxpool_epilogue_matches_torch(case):
    """maxpool2x2_ceil(relu(conv3x3 + bias)) in ONE kernel (modelA_MTL.py:102,116,145) against torch on bf16-rounded
    operands; 5 and 13 images (ragged last tiles: a tile is a whole number of image-row pairs)."""
    from mtldas_b200 import conv2d_nchw
    cin, cout, H, W = case
    for n in (5, 13):
        g = torch.Generator().manual_seed(cin * 7 + n)
        x = torch.randn(n, cin, H, W, generator=g)
        w = torch.randn(cout, cin, 3, 3, generator=g) / (3.0 * cin ** 0.5)
        b = torch.randn(cout, generator=g) * 0.3
        xr, wr = x.bfloat16().float(), w.bfloat16().float()
        y = torch.relu(F.conv2d(xr.double(), wr.double(), b.double(), padding=1))
        ref = F.max_pool2d(y, kernel_size=2, stride=2, ceil_mode=True)
        out = conv2d_nchw(x.cuda(), w.cuda(), b.cuda(), epilogue=5, precision="bf16").cpu().double()
        assert out.shape == ref.shape == (n, cout, (H + 1) // 2, (W + 1) // 2)
        err = float((out - ref).abs().max() / ref.abs().max())
        assert err <= 6e-3, (case, n, err)      # one bf16 rounding of the stored result


def test_umma_descriptor_selftest():
    from mtldas_b200 import umma_selftest
    table = {}
    for n in (16, 32, 64, 128):
        for shift in (0, 1, 5, 85):
            table[(n, shift)] = umma_selftest(n, shift, 0)
    print("umma selftest max-abs-err:", table)
    assert all(v == 0.0 for v in table.values()), table


@pytest.mark.parametrize("precision", ["fp32", "bf16"])
@pytest.mark.parametrize("case", CONV_CASES, ids=lambda c: "c%d-%d_k%d_s%d_%dx%d_e%d" % c[:7])
def test_conv_matches_torch(case, precision):
    from mtldas_b200 import conv2d_nchw
    cin, cout, k, stride, H, W, epi, cin2, aux_s2 = case
    n = 5  # 5 images: several CTA tiles and a ragged last tile
    g = torch.Generator().manual_seed(cin * 131 + cout * 7 + k + stride + epi)
    x = torch.randn(n, cin, H, W, generator=g)
    x2 = torch.randn(n, cin2, H, W, generator=g) if cin2 else None
    w = torch.randn(cout, cin + cin2, k, k, generator=g) / np.sqrt((cin + cin2) * k * k)
    b = torch.randn(cout, generator=g) * 0.1
    ho, wo = (H, W) if stride == 1 else ((H + 1) // 2, (W + 1) // 2)
    aux = torch.randn(n, cout, ho, wo, generator=g) if epi in (2, 3) else None
    if precision == "bf16":   # the kernel's operand rounding points
        x, w = O.bf16_round(x), O.bf16_round(w)
        x2 = O.bf16_round(x2) if x2 is not None else None
        aux = O.bf16_round(aux) if aux is not None else None
    ref = _ref_conv(x, w, b, aux, x2, stride, epi, k)
    dev = "cuda:0"
    got = conv2d_nchw(x.to(dev), w.to(dev), b.to(dev), aux.to(dev) if aux is not None else None,
                      x2.to(dev) if x2 is not None else None, stride=stride, epilogue=epi, precision=precision,
                      aux_phase_split=bool(aux_s2)).cpu().double()
    assert got.shape == ref.shape
    scale = float(ref.abs().max())
    err = float((got - ref).abs().max())
    # fp32 path: true fp32 FMAs -> 1e-4 relative (north_star); bf16 path: exact operands, fp32 accumulate,
    # result rounded to bf16 on store -> 2^-8 relative
    tol = 1e-4 if precision == "fp32" else 6e-3
    assert err <= tol * scale, (err, scale)


@pytest.mark.parametrize("kind", ["MTL", "single_distance", "single_event"])
@pytest.mark.parametrize("precision", ["fp32", "bf16"])
def test_forward_eval_matches_oracle(kind, precision):
    from mtldas_b200 import EvalPlan
    n = 11   # micro-batch 4 -> passes of 4,4,3 (ragged last pass)
    sd = O.synth_state_dict(kind, seed=0)
    x, _, _ = O.synth_patches(n, seed=4321)
    ref = O.forward(sd, x, kind, training=False, dtype=torch.float64)
    plan = EvalPlan(kind, 4, precision, "cuda:0")
    plan.pack({k: v.cuda() for k, v in sd.items()})
    outs = plan.forward(x.cuda())
    torch.cuda.synchronize()
    tol = 1e-4 if precision == "fp32" else 1e-2   # north_star tolerances (relative, on log-probs)
    for o, r in zip(outs, ref):
        o = o.cpu().double()
        assert o.shape == r.shape
        err = float((o - r).abs().max() / r.abs().max())
        assert err <= tol, (kind, precision, err)
        if precision == "fp32":
            assert torch.equal(o.argmax(1), r.argmax(1))


def test_intermediate_tensors_fp32():
    """Localises a failure: shared features of the last pass vs the oracle's feature maps."""
    from mtldas_b200 import EvalPlan
    n = 3
    sd = O.synth_state_dict("MTL", seed=0)
    x, _, _ = O.synth_patches(n, seed=5)
    _, feats = O.forward(sd, x, "MTL", training=False, dtype=torch.float64, return_feats=True)
    plan = EvalPlan("MTL", 4, "fp32", "cuda:0")
    plan.pack({k: v.cuda() for k, v in sd.items()})
    plan.forward(x.cuda())
    for name in ["stem"] + [f"sb{i}" for i in range(8)]:
        got = plan.read_tensor(name, n).cpu().double()
        ref = feats[name]
        err = float((got - ref).abs().max() / ref.abs().max())
        assert err < 1e-4, (name, err)


def test_dropin_module_and_host_path():
    from model.modelA_MTL import MTL_Net
    n = 9
    sd = O.synth_state_dict("MTL", seed=0)
    x, _, _ = O.synth_patches(n, seed=77)
    ref = O.forward(sd, x, "MTL", training=False, dtype=torch.float64)
    m = MTL_Net()
    m.precision, m.micro_batch = "fp32", 4
    m.load_state_dict(sd, strict=True)
    m = m.cuda().eval()
    out1, out2 = m(x.cuda())
    assert out1.shape == (n, 16) and out2.shape == (n, 2)
    for o, r in zip((out1, out2), ref):
        assert float((o.cpu().double() - r).abs().max() / r.abs().max()) < 1e-4
    # host-buffer entry point == device entry point
    plan = m._eval_plan(x.cuda().device, n)
    h1, h2 = plan.forward_host(x.pin_memory())
    assert torch.allclose(h1, out1.cpu(), atol=1e-6) and torch.allclose(h2, out2.cpu(), atol=1e-6)
    # a parameter update invalidates the packed weights
    with torch.no_grad():
        m.conv1[0].weight.mul_(1.5)
    o1b, _ = m(x.cuda())
    assert float((o1b - out1).abs().max()) > 1e-4


def test_argmax_agreement_bf16_batch():
    """north_star: bf16 path argmax agreement >= 99.9 % (near-ties excluded as in SURVEY App. E.4)."""
    from mtldas_b200 import EvalPlan
    n = 512
    sd = O.synth_state_dict("MTL", seed=0)
    x, _, _ = O.synth_patches(n, seed=2024)
    ref = O.forward(sd, x, "MTL", training=False, dtype=torch.float64)
    plan = EvalPlan("MTL", 128, "bf16", "cuda:0")
    plan.pack({k: v.cuda() for k, v in sd.items()})
    outs = [o.cpu().double() for o in plan.forward(x.cuda())]
    for o, r in zip(outs, ref):
        maxerr = float((o - r).abs().max())
        top2 = r.topk(2, dim=1).values
        margin = top2[:, 0] - top2[:, 1]
        clear = margin > 4 * maxerr
        agree = (o.argmax(1) == r.argmax(1))
        assert float(agree[clear].double().mean()) >= 0.999
        print("bf16 argmax raw agreement", float(agree.double().mean()), "clear", int(clear.sum()), "maxerr", maxerr)


def test_streaming_inference_windows_match_oracle():
    """Large-array streaming inference: every (window, time block) of a synthetic array, sharded over
    two emulated ranks with no exchange, equals the oracle on the explicitly cut patches."""
    from mtldas_b200.stream_infer import StreamingInference
    sd = O.synth_state_dict("MTL", seed=0)
    g = torch.Generator().manual_seed(3)
    arr = torch.randn(160, 750, generator=g)          # 160 fibre channels x 3 time blocks
    stride = 20
    eng = StreamingInference(sd, "MTL", "fp32", micro_batch=4, stride=stride, device="cuda:0")
    got0, got1, starts_all = [], [], []
    for rank in range(2):
        starts, (lo, hi) = eng.windows_of_rank(arr.shape[0], rank, 2)
        starts_all += starts
        slab = arr[lo:hi].cuda()                       # the rank holds only its channel slab
        o0, o1 = eng.run(slab, first_channel=0, n_windows=len(starts))
        got0.append(o0.cpu()); got1.append(o1.cpu())
    got0, got1 = torch.cat(got0), torch.cat(got1)
    assert starts_all == list(range(0, 160 - 100 + 1, stride))
    patches = torch.stack([arr[c:c + 100, tb * 250:(tb + 1) * 250] for c in starts_all for tb in range(3)]).unsqueeze(1)
    ref = O.forward(sd, patches, "MTL", training=False, dtype=torch.float64)
    for got, r in zip((got0, got1), ref):
        r = r.view(len(starts_all), 3, -1)
        assert got.shape == r.shape
        assert float((got.double() - r).abs().max() / r.abs().max()) <= 1e-4


def test_tensor_core_stem_matches_oracle_and_cuda_core_stem(monkeypatch):
    """The tcgen05 stem (one banded GEMM per output row, stem_umma.cu) against the fp64 oracle's stem feature map and
    against the CUDA-core stem on the same inputs; also the unaligned-patch fallback (bulk copies need 16-byte
    aligned patches: a tensor view at a 4-byte offset must take the CUDA-core kernel and give the same numbers)."""
    from mtldas_b200 import EvalPlan
    n = 7                                   # 7 images x 3 units: partial last unit (10 rows) and a ragged grid
    sd = O.synth_state_dict("MTL", seed=0)
    x, _, _ = O.synth_patches(n, seed=11)
    _, feats = O.forward(sd, x, "MTL", training=False, dtype=torch.float64, return_feats=True)
    ref = feats["stem"]
    got = {}
    for flag in ("1", "0"):
        monkeypatch.setenv("MTLDAS_UMMA_STEM", flag)
        plan = EvalPlan("MTL", 8, "bf16", "cuda:0")
        plan.pack({k: v.cuda() for k, v in sd.items()})
        plan.forward(x.cuda())
        got[flag] = plan.read_tensor("stem", n).cpu().double()
        err = float((got[flag] - ref).abs().max() / ref.abs().max())
        assert err < 1e-2, (flag, err)       # bf16 operands (flag 1) / bf16 store (both): north_star bf16 tolerance
    # the two kernels differ only by the bf16 rounding of the patch and the weights
    assert float((got["1"] - got["0"]).abs().max() / ref.abs().max()) < 1e-2
    monkeypatch.setenv("MTLDAS_UMMA_STEM", "1")
    plan = EvalPlan("MTL", 8, "bf16", "cuda:0")
    plan.pack({k: v.cuda() for k, v in sd.items()})
    big = torch.empty(n * 25000 + 1, device="cuda")
    xu = big[1:].view(n, 1, 100, 250)        # 4-byte offset: not 16-byte aligned
    xu.copy_(x)
    assert xu.data_ptr() % 16 != 0
    plan.forward(xu)
    fallback = plan.read_tensor("stem", n).cpu().double()
    assert torch.equal(fallback, got["0"])


def test_eval_is_batching_invariant_at_full_size():
    """Size-independent property at BASELINE.json's full inference size: eval-mode results do not depend on how the
    patches are batched (4100 > the 4096-patch plan, so the module runs two passes; the same patches as three ragged
    calls must give bit-identical log-probs on the bf16 path)."""
    from model.modelA_MTL import MTL_Net
    m = MTL_Net()
    m.load_state_dict(O.synth_state_dict("MTL", seed=0), strict=True)
    m = m.cuda().eval()
    base, _, _ = O.synth_patches(64, seed=3)
    g = torch.Generator().manual_seed(5)
    x = (base.repeat(65, 1, 1, 1)[:4100] + 0.1 * torch.randn(4100, 1, 100, 250, generator=g)).cuda()
    with torch.no_grad():
        whole = [o.clone() for o in m(x)]
        parts = [[o.clone() for o in m(x[a:b])] for a, b in ((0, 1), (1, 1000), (1000, 4100))]
    for t in range(2):
        cat = torch.cat([p[t] for p in parts])
        assert not bool(torch.isnan(whole[t]).any())
        assert torch.equal(whole[t], cat)
        assert float((whole[t].exp().sum(1) - 1).abs().max()) < 1e-3    # rows are log-probabilities
