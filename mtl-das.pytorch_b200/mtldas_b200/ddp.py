"""Data-parallel plumbing for the MTL-DAS hot path (one process per GPU, torch.distributed).

Training: the batch is sharded across ranks (BatchNorm statistics stay per replica, as torch DDP does without
SyncBN), every rank runs mtldas_train_forward/_backward on its shard and ONE all-reduce (sum, then 1/world) of the
flat fp32 gradient buffer (1 136 224 elements = 4.54 MB for model A) follows per step -- NCCL over NVLink on B200,
gloo in the CPU tests.  Inference / streaming: the patches (fibre-channel windows) are sharded with no collective.
The reference has no distributed code at all; this module is what replaces "run train.py on
one GPU" when 8 are available.
"""
from typing import List, Sequence, Tuple

import torch
import torch.distributed as dist


def shard_range(n_items: int, rank: int, world: int) -> Tuple[int, int]:
    """Contiguous, balanced [lo, hi) slice of n_items for `rank` (first n_items % world ranks get one more)."""
    if world < 1 or not (0 <= rank < world) or n_items < 0:
        raise ValueError("bad shard arguments")
    base, extra = divmod(n_items, world)
    lo = rank * base + min(rank, extra)
    return lo, lo + base + (1 if rank < extra else 0)


def window_starts(n_channels: int, window: int = 100, stride: int = 10) -> List[int]:
    """First fibre channel of every window of a streaming array [n_channels, T] (full windows only)."""
    if n_channels < window:
        return []
    return list(range(0, n_channels - window + 1, stride))


def shard_windows(n_channels: int, rank: int, world: int, window: int = 100, stride: int = 10):
    """Windows owned by `rank` and the channel slab [c_lo, c_hi) it must hold (its windows plus their halo)."""
    starts = window_starts(n_channels, window, stride)
    lo, hi = shard_range(len(starts), rank, world)
    mine = starts[lo:hi]
    if not mine:
        return [], (0, 0)
    return mine, (mine[0], mine[-1] + window)


def allreduce_mean_(flat: torch.Tensor, group=None) -> torch.Tensor:
    """In-place mean over ranks of the flat gradient buffer: one collective per step."""
    if dist.is_available() and dist.is_initialized() and dist.get_world_size(group) > 1:
        if flat.is_cuda and dist.get_backend(group) == "nccl":
            dist.all_reduce(flat, op=dist.ReduceOp.AVG, group=group)   # the division happens inside the collective
        else:
            dist.all_reduce(flat, op=dist.ReduceOp.SUM, group=group)   # gloo has no AVG
            flat.div_(dist.get_world_size(group))
    return flat


def unflatten_like(flat: torch.Tensor, params: Sequence[torch.Tensor]) -> List[torch.Tensor]:
    """Views of the flat buffer with the shapes of `params` (named_parameters order)."""
    out, off = [], 0
    for p in params:
        n = p.numel()
        out.append(flat[off:off + n].view(p.shape))
        off += n
    if off != flat.numel():
        raise ValueError(f"flat buffer has {flat.numel()} elements, parameters need {off}")
    return out


def flatten_parameters_(module, out: torch.Tensor = None) -> torch.nn.Parameter:
    """Re-home every parameter of `module` into ONE flat fp32 buffer (named_parameters order, the order of the flat
    gradient) and return that buffer as a Parameter.  The module's parameters become views of it, so the state_dict,
    the packing kernels and an optimizer stepping the flat tensor all see the same storage; the elementwise optimizers
    of the reference (Adam with L2 weight decay, utils.py:133-134) then update all 126 tensors in one fused launch
    instead of a multi-tensor sweep."""
    params = [p for _, p in module.named_parameters()]
    if not params:
        raise ValueError("module has no parameters")
    flat = torch.cat([p.detach().reshape(-1).float() for p in params]).contiguous()
    if out is not None:          # caller-provided storage (e.g. a symmetric-memory allocation, possibly padded)
        if out.dtype != torch.float32 or out.numel() < flat.numel() or out.device != flat.device:
            raise ValueError("flatten_parameters_: `out` must be an fp32 buffer on the parameters' device, large enough")
        out.zero_()
        out[:flat.numel()].copy_(flat)
        flat = out
    off = 0
    for p in params:
        n = p.numel()
        p.data = flat[off:off + n].view(p.shape)
        off += n
    return torch.nn.Parameter(flat, requires_grad=True)


def broadcast_module_(module, group=None, src=0):
    """Make every rank start from rank `src`'s parameters AND buffers (BatchNorm running statistics, counters), as torch
    DDP does at construction: the reference builds its model with `MTL_Net()` (utils.py:87, per-process random init), so
    without this the replicas would average gradients of different weights for ever."""
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size(group) < 2:
        return
    with torch.no_grad():
        for t in list(module.parameters()) + list(module.buffers()):
            dist.broadcast(t.data, src=src, group=group)


def replica_checksum(module, group=None):
    """(max over ranks) - (min over ranks) of a checksum of all parameters: 0.0 when the replicas are identical.
    One small all-reduce; used by bench.py and the multi-GPU tests after the timed steps."""
    with torch.no_grad():
        ps = [p.detach().double().reshape(-1) for p in module.parameters()]
        flat = torch.cat(ps)
        w = torch.arange(1, flat.numel() + 1, dtype=torch.float64, device=flat.device) % 1021.0 + 1.0
        c = torch.stack(((flat * w).sum(), flat.abs().sum()))
    if dist.is_available() and dist.is_initialized() and dist.get_world_size(group) > 1:
        hi, lo = c.clone(), c.clone()
        dist.all_reduce(hi, op=dist.ReduceOp.MAX, group=group)
        dist.all_reduce(lo, op=dist.ReduceOp.MIN, group=group)
        return float((hi - lo).abs().max()), [float(v) for v in c]
    return 0.0, [float(v) for v in c]


class DataParallelTrainer:
    """One training step per call: shard -> forward/backward on the local plan -> one all-reduce -> optimizer.

    optimizer: a torch optimizer over module.parameters() (the reference's construction), or None together with
    `flat_adam=dict(lr=..., weight_decay=...)` to flatten the parameters and step ONE Adam launch over the flat buffer
    (mtldas_adam_step on a B200; torch.optim.Adam on the CPU, where only the host logic is exercised).
    Construction broadcasts rank 0's parameters and buffers to every rank of `group`."""

    def __init__(self, module, optimizer=None, group=None, flat_adam=None, broadcast=True, fused_collective="auto"):
        self.module, self.optimizer, self.group = module, optimizer, group
        self._flat_param = None
        self._flat = None
        self._fused = None            # state of the one-kernel reduce + Adam + broadcast (multimem) path
        self.collective = "none"
        if broadcast:
            broadcast_module_(module, group)
        if optimizer is None:
            if flat_adam is None:
                raise ValueError("give an optimizer or flat_adam=dict(lr=..., weight_decay=...)")
            kw = dict(flat_adam)
            first = next(module.parameters())
            multi = dist.is_available() and dist.is_initialized() and dist.get_world_size(group) > 1
            if first.is_cuda and multi and fused_collective in ("auto", True):
                self._setup_fused(module, kw, group, required=fused_collective is True)
            if self._fused is None:
                self._flat_param = flatten_parameters_(module)
                if self._flat_param.is_cuda:
                    from .engine import FlatAdam
                    self.optimizer = FlatAdam(self._flat_param, **kw)
                else:
                    self.optimizer = torch.optim.Adam([self._flat_param], **kw)
                self.collective = "nccl all-reduce (AVG) + mtldas_adam_step" if multi and first.is_cuda else \
                    ("all-reduce + Adam" if multi else "none (single rank)")
        elif dist.is_available() and dist.is_initialized() and dist.get_world_size(group) > 1:
            self.collective = "all-reduce (mean) + caller's optimizer"
        self.params = [p for _, p in module.named_parameters()]
        self._loss_bufs = None
        self._state = None

    def _setup_fused(self, module, kw, group, required):
        """Flat parameter and gradient buffers in symmetric memory with an NVSwitch multicast mapping, so that ONE kernel
        (mtldas_reduce_adam_multimem) sums the gradient slices in the switch, applies Adam and broadcasts the updated
        slice to every replica.  Falls back to NCCL all-reduce + mtldas_adam_step when the mapping is not available."""
        try:
            import torch.distributed._symmetric_memory as symm
            world, rank = dist.get_world_size(group), dist.get_rank(group)
            pg = group if group is not None else dist.group.WORLD
            gname = pg.group_name
            dev = next(module.parameters()).device
            n = sum(p.numel() for p in module.parameters())
            unit = 4 * world
            n_pad = (n + unit - 1) // unit * unit
            pbuf = symm.empty(n_pad, dtype=torch.float32, device=dev)
            gbuf = symm.empty(n_pad, dtype=torch.float32, device=dev)
            hp, hg = symm.rendezvous(pbuf, gname), symm.rendezvous(gbuf, gname)
            if not hp.multicast_ptr or not hg.multicast_ptr:
                raise RuntimeError("no multicast mapping for symmetric memory on this system")
            gbuf.zero_()
            self._flat_param = flatten_parameters_(module, out=pbuf)
            self._flat = gbuf[:n]
            self._flat_param.grad = None
            ns = n_pad // world
            self._fused = dict(hp=hp, hg=hg, pbuf=pbuf, gbuf=gbuf, n=n, n_pad=n_pad, rank=rank, world=world, step=0,
                               m=torch.zeros(ns, dtype=torch.float32, device=dev), v=torch.zeros(ns, dtype=torch.float32, device=dev),
                               lr=float(kw.get("lr", 1e-3)), betas=tuple(kw.get("betas", (0.9, 0.999))),
                               eps=float(kw.get("eps", 1e-8)), weight_decay=float(kw.get("weight_decay", 0.0)))
            hp.barrier(channel=0)
            self.collective = "mtldas_reduce_adam_multimem (multimem.ld_reduce + Adam + multimem.st, one kernel)"
        except Exception as e:          # pragma: no cover - depends on the machine
            self._fused, self._flat = None, None
            if required:
                raise
            self.fused_unavailable = repr(e)[:300]

    @property
    def lr(self):
        return self._fused["lr"] if self._fused else getattr(self.optimizer, "lr", None)

    @lr.setter
    def lr(self, value):
        if self._fused:
            self._fused["lr"] = float(value)
        elif hasattr(self.optimizer, "lr"):
            self.optimizer.lr = float(value)
        else:
            for g in self.optimizer.param_groups:
                g["lr"] = float(value)

    def _fused_step(self, device):
        import ctypes
        from ._lib import lib, check
        from .engine import _stream_ptr
        f = self._fused
        f["step"] += 1
        f["hg"].barrier(channel=0)                     # every rank's gradient is written
        with torch.cuda.device(device):
            check(lib.mtldas_reduce_adam_multimem(
                ctypes.c_void_p(f["hg"].multicast_ptr), ctypes.c_void_p(f["hp"].multicast_ptr),
                ctypes.c_void_p(f["pbuf"].data_ptr()), ctypes.c_void_p(f["m"].data_ptr()), ctypes.c_void_p(f["v"].data_ptr()),
                f["n_pad"], f["rank"], f["world"], f["step"], f["lr"], f["betas"][0], f["betas"][1], f["eps"],
                f["weight_decay"], _stream_ptr(device)), "reduce_adam_multimem")
        f["hp"].barrier(channel=1)                     # every updated slice has landed in every replica

    def step(self, x_local: torch.Tensor, labels_local: Sequence[torch.Tensor]):
        """x_local / labels_local: this rank's shard. Returns the local loss (python float is NOT taken here)."""
        from .engine import nll_loss
        m = self.module
        plan = m._train_plan(x_local.device, x_local.shape[0])
        if self._state is None:   # parameter / buffer tensors are updated in place, their identities are stable
            self._state = [v.detach() for v in m.state_dict(keep_vars=True).values()]
        outs = plan.forward(self._state, x_local)
        # NLLLoss (mean) per task, summed, and its gradient: one fused launch (utils.py:139,361,367)
        self._loss_bufs = nll_loss(list(outs), [y.long().contiguous() for y in labels_local], out=self._loss_bufs)
        loss, douts = self._loss_bufs
        if self._flat is None or self._flat.device != x_local.device:
            self._flat = torch.empty(plan.n_params, dtype=torch.float32, device=x_local.device)
            if self._flat_param is not None:
                self._flat_param.grad = self._flat  # one tensor, one optimizer launch
            else:
                for p, g in zip(self.params, unflatten_like(self._flat, self.params)):
                    p.grad = g                     # views of the flat buffer: bound once, refilled every step
        plan.backward(douts, self._flat)
        if self._fused is not None:
            self._fused_step(x_local.device)       # all-reduce + Adam + broadcast: one kernel between two barriers
        else:
            allreduce_mean_(self._flat, self.group)
            self.optimizer.step()
        # parameters and BatchNorm buffers changed under the module (in-place kernels do not bump torch's version
        # counters): the eval-mode weight image packed earlier is stale
        if hasattr(m, "invalidate_packed"):
            m.invalidate_packed()
        return loss
