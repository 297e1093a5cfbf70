"""GPU tests of the pieces around the hot path that round 2 added or hardened (all through the C ABI):
Adam over the flat buffer, validation metrics on the device, eval-after-training repacking, the stale-forward guard,
bf16 host upload, the reference trainer's call sequence on the drop-in module, and the 2-GPU NCCL training step."""
import os
import sys

import pytest
import torch

from oracle import mtl_oracle as O

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_adam_kernel_matches_torch_adam():
    """mtldas_adam_step vs torch.optim.Adam(lr 1e-3/1.5, weight_decay 1e-5) (utils.py:133-134,247,374), three steps on a
    buffer of model A's size + 3 (ragged tail), against torch's fp32 Adam on the GPU (<= 2e-6 relative: same formula,
    different rounding of the divisions) and an fp64 restatement (<= 1e-6)."""
    from mtldas_b200 import FlatAdam
    n = 1136224 + 3
    g = torch.Generator().manual_seed(3)
    p0 = torch.randn(n, generator=g) * 0.1
    grads = [torch.randn(n, generator=g) * (10.0 ** -k) for k in range(3)]
    lr, wd = 1e-3 / 1.5, 1e-5
    mine = torch.nn.Parameter(p0.clone().cuda())
    opt = FlatAdam(mine, lr=lr, weight_decay=wd)
    ref = torch.nn.Parameter(p0.clone().cuda())
    ropt = torch.optim.Adam([ref], lr=lr, weight_decay=wd)
    p64, m64, v64 = p0.double(), torch.zeros(n, dtype=torch.float64), torch.zeros(n, dtype=torch.float64)
    for t, gr in enumerate(grads, start=1):
        opt.step(gr.cuda())
        ref.grad = gr.cuda()
        ropt.step()
        gd = gr.double() + wd * p64
        m64 = 0.9 * m64 + 0.1 * gd
        v64 = 0.999 * v64 + 0.001 * gd * gd
        p64 = p64 - (lr / (1 - 0.9 ** t)) * m64 / (v64.sqrt() / (1 - 0.999 ** t) ** 0.5 + 1e-8)
    torch.cuda.synchronize()
    a = mine.detach().cpu().double()
    e_torch = float((a - ref.detach().cpu().double()).abs().max() / p64.abs().max())
    e_64 = float((a - p64).abs().max() / p64.abs().max())
    print("adam kernel: max rel error vs torch fp32 Adam", e_torch, "vs fp64 restatement", e_64)
    assert e_torch <= 2e-6 and e_64 <= 1e-6
    assert opt.step_count == 3 and float(opt.exp_avg_sq.min()) >= 0.0


def test_eval_metrics_kernel_matches_torch_bookkeeping():
    """utils.py:274-291,313 on the device: NLL sums, torch.max(out,1)[1] (first maximum wins), correct counts, confusion
    matrices (row = label, column = prediction), accumulated over two batches; integer results are exact."""
    from mtldas_b200 import EvalMetrics
    g = torch.Generator().manual_seed(5)
    met = EvalMetrics((16, 2), "cuda:0")
    conf = [torch.zeros(16, 16, dtype=torch.int64), torch.zeros(2, 2, dtype=torch.int64)]
    correct, seen, nll = [0, 0], 0, [0.0, 0.0]
    for n in (1000, 37):
        lp = [torch.log_softmax(torch.randn(n, 16, generator=g), 1), torch.log_softmax(torch.randn(n, 2, generator=g), 1)]
        lp[0][::7, 3] = lp[0][::7].max(1).values       # exact ties: the lower index must win
        lp[0][::7, 9] = lp[0][::7, 3]
        ys = [torch.randint(0, 16, (n,), generator=g), torch.randint(0, 2, (n,), generator=g)]
        preds = met.update([t.cuda().contiguous() for t in lp], [y.cuda() for y in ys], want_pred=True)
        for t in range(2):
            want = torch.max(lp[t], 1)[1]
            assert torch.equal(preds[t].cpu(), want), t
            correct[t] += int((want == ys[t]).sum())
            nll[t] += float(-lp[t].double().gather(1, ys[t].view(-1, 1)).sum())
            conf[t].index_put_((ys[t], want), torch.ones(n, dtype=torch.int64), accumulate=True)
        seen += n
    r = met.result()
    assert r["seen"] == seen and r["bad_labels"] == 0
    for t in range(2):
        assert torch.equal(r["confusion"][t], conf[t]), t
        assert abs(r["accuracy"][t] - correct[t] / seen) < 1e-12
        assert abs(r["nll_mean"][t] - nll[t] / seen) <= 1e-5 * max(1.0, abs(nll[t] / seen))
    met.reset()
    met.update([torch.zeros(4, 16, device="cuda"), torch.zeros(4, 2, device="cuda")],
               [torch.full((4,), 16, device="cuda"), torch.zeros(4, dtype=torch.int64, device="cuda")])
    assert met.result()["bad_labels"] == 4


def test_eval_after_flat_adam_training_sees_updated_weights_and_statistics():
    """eval, N steps of the flat-buffer trainer, eval again -- the second eval must use the updated
    parameters AND BatchNorm running statistics (nothing bumps torch's version counters on that path)."""
    from model.modelA_MTL import MTL_Net
    from mtldas_b200.ddp import DataParallelTrainer
    sd = O.synth_state_dict("MTL", seed=0)
    m = MTL_Net()
    m.precision = "fp32"
    m.load_state_dict(sd, strict=True)
    m = m.cuda()
    xe, _, _ = O.synth_patches(6, seed=41)
    m.eval()
    first = [o.clone() for o in m(xe.cuda())]
    tr = DataParallelTrainer(m, flat_adam=dict(lr=1e-2, weight_decay=1e-5))
    m.train()
    for step in range(3):
        x, yd, ye = O.synth_patches(8, seed=50 + step)
        tr.step(x.cuda(), [yd.cuda(), ye.cuda()])
    m.eval()
    second = m(xe.cuda())
    now = {k: v.detach().cpu() for k, v in m.state_dict().items()}
    assert int(now["conv1.1.num_batches_tracked"]) == int(sd["conv1.1.num_batches_tracked"]) + 3
    want = O.forward(now, xe, "MTL", training=False, dtype=torch.float64)
    for o, w, f in zip(second, want, first):
        assert float((o.cpu().double() - w).abs().max() / w.abs().max()) <= 1e-4
        assert float((o - f).abs().max()) > 1e-3          # and it did change


def test_backward_of_a_stale_forward_raises():
    from model.modelA_MTL import MTL_Net
    from mtldas_b200 import MtldasError
    m = MTL_Net()
    m.precision = "fp32"
    m.load_state_dict(O.synth_state_dict("MTL", seed=0), strict=True)
    m = m.cuda().train()
    x, yd, ye = O.synth_patches(4, seed=61)
    crit = torch.nn.NLLLoss()
    o1, o2 = m(x.cuda())
    loss_a = crit(o1, yd.cuda()) + crit(o2, ye.cuda())
    p1, p2 = m(x.cuda())                       # a second train-mode forward overwrites the plan's activations
    loss_b = crit(p1, yd.cuda()) + crit(p2, ye.cuda())
    with pytest.raises(MtldasError):
        loss_a.backward()
    loss_b.backward()                          # the most recent forward is fine
    assert m.conv1[0].weight.grad is not None


def test_host_entry_point_bf16_upload_is_bit_identical():
    """mtldas_plan_forward_eval_host_bf16: bf16 host patches (half the PCIe bytes) give exactly the log-probs of the fp32
    upload of the same (bf16-representable) values; ragged chunking included."""
    from mtldas_b200 import EvalPlan
    n = 1100                                             # host chunks of 512, 512, 76
    sd = O.synth_state_dict("MTL", seed=0)
    x, _, _ = O.synth_patches(64, seed=71)
    x = x.repeat(18, 1, 1, 1)[:n].contiguous()
    xh = x.to(torch.bfloat16)
    xf = xh.float()
    plan = EvalPlan("MTL", 512, "bf16", "cuda:0")
    plan.pack({k: v.cuda() for k, v in sd.items()})
    a = [o.clone() for o in plan.forward_host(xf.pin_memory())]
    b = [o.clone() for o in plan.forward_host(xh.pin_memory())]
    c = [o.cpu() for o in plan.forward(x.cuda())]        # unrounded fp32 patches: the stem rounds to the same bf16
    for u, v, w in zip(a, b, c):
        assert torch.equal(u, v)
        assert torch.equal(u, w)
    fp32_plan = EvalPlan("MTL", 8, "fp32", "cuda:0")
    fp32_plan.pack({k: v.cuda() for k, v in sd.items()})
    from mtldas_b200 import MtldasError
    with pytest.raises(MtldasError):
        fp32_plan.forward_host(xh[:4].contiguous().pin_memory())


def test_reference_trainer_call_sequence_on_the_dropin_module(tmp_path):
    """The call sequence of the unmodified trainer (utils.py:226-403) replayed on the drop-in module: eval pass with
    grad ENABLED (utils.py:249-272 has no torch.no_grad), metrics, train loop with
    optim.Adam(model.parameters(), 1e-3, weight_decay=1e-5) / NLLLoss / zero_grad-backward-step incl. a partial last
    batch (DataLoader drop_last=False), eval again, torch.save(state_dict) -> load_state_dict(strict=True) into a fresh
    module (what test.py does, utils.py:122-123) -> same outputs; the saved file also loads into the ORACLE's key set.
    (The reference tree is absent on the GPU box and the build container has no GPU, so the real train.py cannot run on
    this path anywhere; mtldas_b200.run_reference is what binds it, tests/test_plan_program.py covers that binding.)"""
    from model.modelA_MTL import MTL_Net
    torch.manual_seed(0)
    model = MTL_Net()                                     # reference construction: default init
    model.precision = "fp32"
    model = model.cuda()
    optimizer = torch.optim.Adam(model.parameters(), lr=1e-3, weight_decay=1e-5)
    criterion = torch.nn.NLLLoss()
    xs, yds, yes = O.synth_patches(44, seed=81)           # batches of 16, 16, 12
    batches = [(xs[i:i + 16], yds[i:i + 16], yes[i:i + 16]) for i in range(0, 44, 16)]

    def validate():
        model.eval()
        acc, tot, loss = [0, 0], 0, 0.0
        for x, yd, ye in batches:
            out1, out2 = model(x.cuda())                  # grad mode is on, as in the reference
            loss += float(criterion(out1, yd.cuda().long()).cpu().data) + float(criterion(out2, ye.cuda().long()).cpu().data)
            acc[0] += int((torch.max(out1, 1)[1] == yd.cuda()).sum())
            acc[1] += int((torch.max(out2, 1)[1] == ye.cuda()).sum())
            tot += len(yd)
        return loss / len(batches), acc[0] / tot, acc[1] / tot

    l0, _, _ = validate()
    first_train_loss = last_train_loss = None
    for epoch in range(3):
        model.train(True)
        for x, yd, ye in batches:
            out1, out2 = model(x.cuda())
            loss = criterion(out1, yd.cuda().long()) + criterion(out2, ye.cuda().long())
            optimizer.zero_grad()
            loss.backward()
            optimizer.step()
            v = float(loss.detach())
            first_train_loss = v if first_train_loss is None else first_train_loss
            last_train_loss = v
    l1, a1, a2 = validate()
    print("reference-protocol run: val loss", l0, "->", l1, "train loss", first_train_loss, "->", last_train_loss, "acc", a1, a2)
    assert last_train_loss < first_train_loss and l1 == l1
    path = os.path.join(str(tmp_path), "ckpt.pth")
    torch.save(model.state_dict(), path)
    sd = torch.load(path, map_location="cpu")
    assert list(sd.keys()) == [k for k, _, _ in O.state_dict_keys("MTL")]
    fresh = MTL_Net()
    fresh.precision = "fp32"
    fresh.load_state_dict(sd, strict=True)
    fresh = fresh.cuda().eval()
    model.eval()
    x = batches[0][0]
    want = O.forward(sd, x, "MTL", training=False, dtype=torch.float64)
    for a, b, w in zip(model(x.cuda()), fresh(x.cuda()), want):
        assert torch.equal(a, b)
        assert float((a.cpu().double() - w).abs().max() / w.abs().max()) <= 1e-4


# ---------------------------------------------------------------------------------------------------------------
# two real GPUs, NCCL
# ---------------------------------------------------------------------------------------------------------------
def _nccl_worker(rank, world, port, tmp):
    sys.path[:0] = [ROOT, os.path.join(ROOT, "mtl-das.pytorch_b200")]
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    import torch.distributed as dist
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", rank))
    from model.modelA_MTL import MTL_Net
    from mtldas_b200.ddp import DataParallelTrainer, replica_checksum, shard_range
    from oracle import mtl_oracle as Or
    sd = Or.synth_state_dict("MTL", seed=0)
    torch.manual_seed(7 + rank)                      # different initial weights per rank: construction must broadcast
    m = MTL_Net()
    if rank == 0:
        m.load_state_dict(sd, strict=True)
    m.precision = "fp32"
    m = m.cuda()
    # NCCL path (the fused multimem path reduces inside its kernel and never materialises the mean gradient); lr 0:
    # the step leaves the weights alone
    tr = DataParallelTrainer(m, flat_adam=dict(lr=0.0, weight_decay=0.0), fused_collective=False)
    spread0, _ = replica_checksum(m)
    n = 12
    x, yd, ye = Or.synth_patches(n, seed=91)
    lo, hi = shard_range(n, rank, world)
    m.train()
    tr.step(x[lo:hi].cuda(), [yd[lo:hi].cuda(), ye[lo:hi].cuda()])
    flat = tr._flat.detach().cpu().double()
    # per-shard fp64 oracle gradients, mean over ranks (per-replica BatchNorm statistics)
    names = Or.param_names("MTL")
    want = None
    for r in range(world):
        a, b = shard_range(n, r, world)
        g = Or.loss_and_grads(sd, x[a:b], yd[a:b], ye[a:b], "MTL", dtype=torch.float64)[2]
        f = torch.cat([g[k].reshape(-1) for k in names])
        want = f if want is None else want + f
    want /= world
    err = float((flat - want).norm() / want.norm())
    spread1, sums = replica_checksum(m)
    torch.save({"err": err, "spread0": spread0, "spread1": spread1, "sums": sums}, os.path.join(tmp, f"n{rank}.pt"))
    dist.destroy_process_group()


def test_two_gpu_nccl_training_step_equals_mean_of_shard_oracle_gradients(tmp_path):
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs (gpurun --gpus 2)")
    import torch.multiprocessing as mp
    world, port = 2, 29100 + os.getpid() % 300
    mp.spawn(_nccl_worker, args=(world, port, str(tmp_path)), nprocs=world, join=True)
    res = [torch.load(os.path.join(str(tmp_path), f"n{r}.pt")) for r in range(world)]
    print("2-GPU NCCL step:", res)
    for r in res:
        assert r["spread0"] == 0.0 and r["spread1"] == 0.0, r       # replicas identical after construction and after the step
        assert r["err"] <= 2e-3, r                                   # fp32 path tolerance of test_train_gradients_fp32
    assert res[0]["sums"] == res[1]["sums"]


def test_streaming_windows_read_in_place_equal_gathered_patches():
    """mtldas_plan_forward_eval_windows (the stem cuts the windows out of the array itself) against
    the gather-then-forward path on the same array: bit-identical log-probs.  Three time blocks (blocks alternate between
    byte offsets 0 and 8 inside a 16-byte line), a non-zero first channel, several micro-batch passes, and the oracle on
    explicitly cut patches."""
    from mtldas_b200.stream_infer import StreamingInference
    sd = O.synth_state_dict("MTL", seed=0)
    g = torch.Generator().manual_seed(9)
    n_ch, n_time, stride = 190, 752, 10                 # 752 % 4 == 0, 3 full time blocks + 2 spare samples
    arr = torch.randn(n_ch, n_time, generator=g)
    direct = StreamingInference(sd, "MTL", "bf16", micro_batch=8, stride=stride, device="cuda:0", in_place=True)
    gathered = StreamingInference(sd, "MTL", "bf16", micro_batch=8, stride=stride, device="cuda:0", in_place=False)
    slab = arr.cuda()
    for first, nwin in ((0, None), (30, 5)):
        a = direct.run(slab, first_channel=first, n_windows=nwin)
        b = gathered.run(slab, first_channel=first, n_windows=nwin)
        for u, v in zip(a, b):
            assert u.shape == v.shape and u.shape[1] == 3
            assert torch.equal(u, v)
    starts = list(range(0, n_ch - 100 + 1, stride))
    patches = torch.stack([arr[c:c + 100, tb * 250:(tb + 1) * 250] for c in starts for tb in range(3)]).unsqueeze(1)
    ref = O.forward(sd, patches, "MTL", training=False, dtype=torch.float64)
    got = direct.run(slab)
    for u, r in zip(got, ref):
        r = r.view(len(starts), 3, -1)
        assert float((u.cpu().double() - r).abs().max() / r.abs().max()) <= 1e-2
    # an array whose row pitch is not a multiple of 4 floats is refused by the C entry point (the wrapper then gathers)
    odd = torch.randn(120, 751, generator=g).cuda()
    c = direct.run(odd)
    d = gathered.run(odd)
    assert all(torch.equal(u, v) for u, v in zip(c, d))


def _fused_worker(rank, world, port, tmp):
    sys.path[:0] = [ROOT, os.path.join(ROOT, "mtl-das.pytorch_b200")]
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    import torch.distributed as dist
    torch.cuda.set_device(rank)
    dev = torch.device("cuda", rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)
    from model.modelA_MTL import MTL_Net
    from mtldas_b200.ddp import DataParallelTrainer, replica_checksum, shard_range
    from oracle import mtl_oracle as Or
    sd = Or.synth_state_dict("MTL", seed=0)
    res = {}
    # ---- (1) the kernel alone, on injected per-rank gradients: exact comparison with an fp64 restatement --------
    m = MTL_Net()
    m.load_state_dict(sd, strict=True)
    m.precision = "fp32"
    m = m.cuda().train()
    tr = DataParallelTrainer(m, flat_adam=dict(lr=1e-3, weight_decay=1e-5), fused_collective=True)
    res["fused_collective"] = tr.collective
    n = tr._flat.numel()
    p64 = torch.cat([p.detach().reshape(-1) for p in m.parameters()]).cpu().double()
    m64, v64 = torch.zeros_like(p64), torch.zeros_like(p64)
    for step in range(1, 4):
        gs = [torch.randn(n, generator=torch.Generator().manual_seed(1000 * step + r)) * (10.0 ** -(step + r)) for r in range(world)]
        tr._flat.copy_(gs[rank].to(dev))
        tr._fused_step(dev)
        g = sum(x.double() for x in gs) / world + 1e-5 * p64
        m64 = 0.9 * m64 + 0.1 * g
        v64 = 0.999 * v64 + 0.001 * g * g
        p64 = p64 - (1e-3 / (1 - 0.9 ** step)) * m64 / (v64.sqrt() / (1 - 0.999 ** step) ** 0.5 + 1e-8)
    torch.cuda.synchronize()
    got = torch.cat([p.detach().reshape(-1) for p in m.parameters()]).cpu().double()
    res["kernel_err"] = float((got - p64).abs().max() / p64.abs().max())
    res["kernel_spread"], _ = replica_checksum(m)
    # ---- (2) three real training steps, fused vs NCCL all-reduce + mtldas_adam_step ------------------------------
    finals, init = {}, None
    for mode in ("fused", "nccl"):
        m = MTL_Net()
        m.load_state_dict(sd, strict=True)
        m.precision = "fp32"
        m = m.cuda().train()
        tr = DataParallelTrainer(m, flat_adam=dict(lr=1e-3, weight_decay=1e-5), fused_collective=True if mode == "fused" else False)
        res[mode + "_collective"] = tr.collective
        init = torch.cat([p.detach().reshape(-1) for p in m.parameters()]).cpu().double()
        nb = 8
        for step in range(3):
            x, yd, ye = Or.synth_patches(nb, seed=200 + step)
            lo, hi = shard_range(nb, rank, world)
            tr.step(x[lo:hi].cuda(), [yd[lo:hi].cuda(), ye[lo:hi].cuda()])
        torch.cuda.synchronize()
        res[mode + "_spread"], _ = replica_checksum(m)
        finals[mode] = torch.cat([p.detach().reshape(-1) for p in m.parameters()]).cpu().double()
    # Adam's first steps move every parameter by ~lr * sign(g): elements whose gradient is at noise level (atomics, a
    # flipped ReLU mask) can differ by a whole step between ANY two runs, so the statement is about the update as a whole
    res["update_rel_l2"] = float((finals["fused"] - finals["nccl"]).norm() / (finals["nccl"] - init).norm())
    torch.save(res, os.path.join(tmp, f"f{rank}.pt"))
    dist.destroy_process_group()


def test_two_gpu_fused_multimem_reduce_adam_equals_nccl_allreduce_plus_adam(tmp_path):
    """mtldas_reduce_adam_multimem (gradient all-reduce + Adam + parameter broadcast in ONE kernel over NVSwitch
    multicast): (1) on injected per-rank gradients, three steps against an fp64 restatement of
    mean-gradient Adam (<= 1e-6), replicas bit-identical; (2) three real training steps against the NCCL all-reduce +
    mtldas_adam_step trainer: the parameter updates agree (relative L2 of the update), replicas bit-identical."""
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs (gpurun --gpus 2)")
    import torch.multiprocessing as mp
    world, port = 2, 28800 + os.getpid() % 300
    mp.spawn(_fused_worker, args=(world, port, str(tmp_path)), nprocs=world, join=True)
    res = [torch.load(os.path.join(str(tmp_path), f"f{r}.pt")) for r in range(world)]
    print("fused collective vs NCCL:", res[0])
    for r in res:
        assert "multimem" in r["fused_collective"], r
        assert r["kernel_err"] <= 1e-6 and r["kernel_spread"] == 0.0, r
        assert r["fused_spread"] == 0.0 and r["nccl_spread"] == 0.0, r
        assert r["update_rel_l2"] <= 5e-2, r
