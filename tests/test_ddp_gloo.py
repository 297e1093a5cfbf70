"""world_size-2 gloo tests (CPU) of the data-parallel host logic: sharding, window sharding, and the single
flat-gradient all-reduce whose result must equal the mean of per-shard reference gradients."""
import os
import sys

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_shard_range_and_windows():
    from mtldas_b200.ddp import shard_range, shard_windows, window_starts
    for n in (0, 1, 7, 8, 4096, 4099):
        for world in (1, 2, 3, 8):
            spans = [shard_range(n, r, world) for r in range(world)]
            assert spans[0][0] == 0 and spans[-1][1] == n
            assert all(a[1] == b[0] for a, b in zip(spans, spans[1:]))
            sizes = [hi - lo for lo, hi in spans]
            assert max(sizes) - min(sizes) <= 1
    starts = window_starts(8192, 100, 10)
    assert starts[0] == 0 and starts[-1] + 100 <= 8192 and len(starts) == (8192 - 100) // 10 + 1
    seen = []
    for r in range(8):
        mine, (lo, hi) = shard_windows(8192, r, 8, 100, 10)
        assert lo == mine[0] and hi == mine[-1] + 100          # slab = own windows + (window - stride) halo
        seen += mine
    assert seen == starts                                        # every window exactly once, in order
    assert window_starts(50) == [] and shard_windows(50, 0, 2) == ([], (0, 0))
    with pytest.raises(ValueError):
        shard_range(4, 2, 2)


def _worker(rank, world, port, tmp):
    sys.path[:0] = [ROOT, os.path.join(ROOT, "mtl-das.pytorch_b200")]
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from oracle import mtl_oracle as O
    from mtldas_b200.ddp import allreduce_mean_, shard_range, unflatten_like
    torch.set_num_threads(2)
    kind, n = "single_event", 4
    sd = O.synth_state_dict(kind, seed=0)
    x, yd, ye = O.synth_patches(n, seed=31)
    lo, hi = shard_range(n, rank, world)
    _, _, g, _ = O.loss_and_grads(sd, x[lo:hi], yd[lo:hi], ye[lo:hi], kind)     # per-replica BatchNorm statistics
    names = O.param_names(kind)
    flat = torch.cat([g[k].reshape(-1) for k in names]).contiguous()
    allreduce_mean_(flat)
    views = unflatten_like(flat, [sd[k] for k in names])
    # reference: mean over ranks of per-shard gradients, computed locally
    want = None
    for r in range(world):
        a, b = shard_range(n, r, world)
        _, _, gr, _ = O.loss_and_grads(sd, x[a:b], yd[a:b], ye[a:b], kind)
        fr = torch.cat([gr[k].reshape(-1) for k in names])
        want = fr if want is None else want + fr
    want /= world
    ok = bool(torch.allclose(flat, want, rtol=1e-5, atol=1e-7)) and views[0].shape == sd[names[0]].shape
    torch.save({"ok": ok, "err": float((flat - want).abs().max())}, os.path.join(tmp, f"r{rank}.pt"))
    dist.destroy_process_group()


def test_flat_gradient_allreduce_equals_mean_of_shard_gradients(tmp_path):
    world, port = 2, 29600 + os.getpid() % 300
    mp.spawn(_worker, args=(world, port, str(tmp_path)), nprocs=world, join=True)
    for r in range(world):
        res = torch.load(os.path.join(str(tmp_path), f"r{r}.pt"))
        assert res["ok"], res


def _bcast_worker(rank, world, port, tmp):
    sys.path[:0] = [ROOT, os.path.join(ROOT, "mtl-das.pytorch_b200")]
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from mtldas_b200.ddp import DataParallelTrainer, replica_checksum
    torch.manual_seed(100 + rank)                      # the reference constructs MTL_Net() unseeded: ranks differ
    m = torch.nn.Sequential(torch.nn.Conv2d(1, 4, 3), torch.nn.BatchNorm2d(4), torch.nn.Conv2d(4, 2, 1))
    m[1].running_mean.add_(rank + 1.0)                 # buffers differ too
    before, _ = replica_checksum(m)
    tr = DataParallelTrainer(m, flat_adam=dict(lr=1e-3, weight_decay=1e-5))     # broadcasts rank 0's state
    after, sums = replica_checksum(m)
    rm = m[1].running_mean.clone()
    dist.all_reduce(rm, op=dist.ReduceOp.MAX)
    same_buffers = bool(torch.equal(rm, m[1].running_mean))
    # one identical "gradient" step on every rank keeps the replicas identical
    tr._flat_param.grad = torch.ones_like(tr._flat_param)
    tr.optimizer.step()
    final, _ = replica_checksum(m)
    torch.save({"before": before, "after": after, "final": final, "same_buffers": same_buffers},
               os.path.join(tmp, f"b{rank}.pt"))
    dist.destroy_process_group()


def test_trainer_construction_broadcasts_rank0_state(tmp_path):
    """ranks built from different seeds must hold rank 0's parameters and buffers after construction."""
    world, port = 2, 29300 + os.getpid() % 300
    mp.spawn(_bcast_worker, args=(world, port, str(tmp_path)), nprocs=world, join=True)
    for r in range(world):
        res = torch.load(os.path.join(str(tmp_path), f"b{r}.pt"))
        assert res["before"] > 0 and res["after"] == 0.0 and res["final"] == 0.0 and res["same_buffers"], res


def test_flat_parameter_adam_equals_per_tensor_adam():
    """flatten_parameters_ re-homes the parameters into one buffer (named_parameters order = flat-gradient order);
    Adam with L2 weight decay (utils.py:133-134) is elementwise, so stepping the flat tensor must give bit-identical
    parameters to stepping the separate tensors, and the module must keep seeing the updated values."""
    import torch
    from mtldas_b200.ddp import flatten_parameters_, unflatten_like
    torch.manual_seed(0)

    def make():
        torch.manual_seed(1)
        return torch.nn.Sequential(torch.nn.Conv2d(1, 4, 3, bias=False), torch.nn.BatchNorm2d(4), torch.nn.Conv2d(4, 2, 1))
    a, b = make(), make()
    opt_a = torch.optim.Adam(a.parameters(), lr=1e-3 / 1.5, weight_decay=1e-5)
    flat = flatten_parameters_(b)
    opt_b = torch.optim.Adam([flat], lr=1e-3 / 1.5, weight_decay=1e-5)
    names_a = [n for n, _ in a.named_parameters()]
    assert [n for n, _ in b.named_parameters()] == names_a
    assert flat.numel() == sum(p.numel() for p in a.parameters())
    x = torch.randn(3, 1, 8, 8)
    for step in range(3):
        grads = [torch.randn_like(p) for p in a.parameters()]
        for p, g in zip(a.parameters(), grads):
            p.grad = g.clone()
        flat.grad = torch.cat([g.reshape(-1) for g in grads])
        opt_a.step(); opt_b.step()
        for pa, pb, view in zip(a.parameters(), b.parameters(), unflatten_like(flat.data, list(b.parameters()))):
            assert torch.equal(pa.data, pb.data) and pb.data.data_ptr() == view.data_ptr()
        assert torch.equal(a(x), b(x))
    assert set(b.state_dict().keys()) == set(a.state_dict().keys())
