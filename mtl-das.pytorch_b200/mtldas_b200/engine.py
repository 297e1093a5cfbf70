"""Plan objects over the C ABI: device memory and streams come from torch, compute from the library."""
import ctypes
import json

import torch

from ._lib import lib, check, MtldasError

KINDS = {"MTL": 0, "single_distance": 1, "single_event": 2}
PRECISIONS = {"bf16": 0, "fp32": 1}


def _stream_ptr(device):
    return ctypes.c_void_p(torch.cuda.current_stream(device).cuda_stream)


_OK_DEVICES = set()


def _require_b200(device):
    key = str(device)
    if key in _OK_DEVICES:          # the device query behind mtldas_device_ok costs tens of milliseconds
        return
    if not torch.cuda.is_available():
        raise MtldasError("mtldas_b200 needs a CUDA device (B200, sm_100a); there is no CPU fallback")
    with torch.cuda.device(device):
        if not lib.mtldas_device_ok():
            raise MtldasError("mtldas_b200 kernels are built for sm_100a (B200) only")
    _OK_DEVICES.add(key)


class EvalPlan:
    """Eval-mode forward of MTL_Net / Single_Task_Net (utils.py:249-272) through mtldas_plan_*."""

    def __init__(self, kind="MTL", micro_batch=256, precision="bf16", device=None, need_device=True):
        self.kind, self.precision = kind, precision
        self.micro_batch = int(micro_batch)
        self._h = lib.mtldas_plan_create(KINDS[kind], self.micro_batch, PRECISIONS[precision])
        if not self._h:
            raise MtldasError("plan_create: " + lib.mtldas_last_error().decode())
        self.n_state = lib.mtldas_plan_num_state_entries(self._h)
        self.keys = [lib.mtldas_plan_state_key(self._h, i).decode() for i in range(self.n_state)]
        self.workspace_bytes = lib.mtldas_plan_workspace_bytes(self._h)
        self.classes = {"MTL": (16, 2), "single_distance": (16,), "single_event": (2,)}[kind]
        self.device = None
        self._ws = None
        self._state_keepalive = None
        if need_device:
            self.device = torch.device("cuda", torch.cuda.current_device()) if device is None else torch.device(device)
            _require_b200(self.device)
            with torch.cuda.device(self.device):
                self._ws = torch.empty(self.workspace_bytes + 256, dtype=torch.uint8, device=self.device)
                ptr = (self._ws.data_ptr() + 255) & ~255
                check(lib.mtldas_plan_bind_workspace(self._h, ctypes.c_void_p(ptr), self.workspace_bytes,
                                                     _stream_ptr(self.device)), "bind_workspace")

    def __del__(self):
        h = getattr(self, "_h", None)
        if h:
            lib.mtldas_plan_destroy(h)
            self._h = None

    def describe(self):
        n = lib.mtldas_plan_describe(self._h, None, 0)
        buf = ctypes.create_string_buffer(n)
        lib.mtldas_plan_describe(self._h, buf, n)
        return json.loads(buf.value.decode())

    def pack(self, state_dict):
        """state_dict: {key: CUDA tensor} with the reference's keys (load_state_dict contract)."""
        tensors = []
        for k in self.keys:
            t = state_dict[k]
            if t.device != self.device:
                raise MtldasError(f"state entry {k} is on {t.device}, plan on {self.device}")
            want = torch.int64 if k.endswith("num_batches_tracked") else torch.float32
            if t.dtype != want or not t.is_contiguous():
                t = t.to(want).contiguous()
            tensors.append(t)
        arr = (ctypes.c_void_p * self.n_state)(*[t.data_ptr() for t in tensors])
        with torch.cuda.device(self.device):
            check(lib.mtldas_plan_pack_eval(self._h, arr, self.n_state, _stream_ptr(self.device)), "pack_eval")
        self._state_keepalive = tensors

    def forward(self, x, outs=None):
        if not x.is_cuda:
            raise MtldasError("input must be a CUDA tensor; there is no CPU fallback")
        if x.dim() != 4 or tuple(x.shape[1:]) != (1, 100, 250):
            raise ValueError(f"expected [N,1,100,250], got {tuple(x.shape)}")
        x = x.contiguous().float()
        n = x.shape[0]
        if outs is None:
            outs = [torch.empty((n, c), dtype=torch.float32, device=x.device) for c in self.classes]
        with torch.cuda.device(self.device):
            check(lib.mtldas_plan_forward_eval(self._h, ctypes.c_void_p(x.data_ptr()), n,
                                               ctypes.c_void_p(outs[0].data_ptr()),
                                               ctypes.c_void_p(outs[1].data_ptr()) if len(outs) > 1 else None,
                                               _stream_ptr(self.device)), "forward_eval")
        return tuple(outs)

    def forward_host(self, x_host, outs_host=None):
        """Host buffers in, host buffers out (H2D + forward + D2H + sync inside the call)."""
        n = x_host.shape[0]
        if x_host.is_cuda or not x_host.is_contiguous() or x_host.dtype not in (torch.float32, torch.bfloat16):
            raise MtldasError("forward_host takes a contiguous fp32 or bf16 HOST tensor [n,1,100,250]")
        if outs_host is None:
            outs_host = [torch.empty((n, c), dtype=torch.float32).pin_memory() for c in self.classes]
        fn = lib.mtldas_plan_forward_eval_host_bf16 if x_host.dtype == torch.bfloat16 else lib.mtldas_plan_forward_eval_host
        with torch.cuda.device(self.device):
            check(fn(self._h, ctypes.c_void_p(x_host.data_ptr()), n,
                                                    ctypes.c_void_p(outs_host[0].data_ptr()),
                                                    ctypes.c_void_p(outs_host[1].data_ptr()) if len(outs_host) > 1 else None,
                                                    _stream_ptr(self.device)), "forward_eval_host")
        return tuple(outs_host)

    def profile(self, x):
        """Per-op device milliseconds of one forward (CUDA events around every launch)."""
        n = x.shape[0]
        outs = [torch.empty((n, c), dtype=torch.float32, device=x.device) for c in self.classes]
        nops = len(self.describe()["program"])
        ms = (ctypes.c_float * nops)()
        with torch.cuda.device(self.device):
            check(lib.mtldas_plan_profile_eval(self._h, ctypes.c_void_p(x.data_ptr()), n,
                                               ctypes.c_void_p(outs[0].data_ptr()),
                                               ctypes.c_void_p(outs[1].data_ptr()) if len(outs) > 1 else None,
                                               _stream_ptr(self.device), ms, nops), "profile_eval")
        return list(ms)

    @property
    def last_launch_count(self):
        return lib.mtldas_plan_last_launch_count(self._h)

    def read_tensor(self, name, nb):
        d = self.describe()
        t = next(t for t in d["tensors"] if t["name"] == name)
        out = torch.empty((nb, t["C"], t["H"], t["W"]), dtype=torch.float32, device=self.device)
        with torch.cuda.device(self.device):
            check(lib.mtldas_plan_read_tensor(self._h, name.encode(), nb, ctypes.c_void_p(out.data_ptr()),
                                              _stream_ptr(self.device)), "read_tensor")
        return out


class TrainPlan:
    """Train-mode forward (batch-statistics BatchNorm) + backward through mtldas_train_* (utils.py:344-373)."""

    def __init__(self, kind="MTL", batch=32, precision="bf16", device=None, need_device=True):
        self.kind, self.precision, self.batch = kind, precision, int(batch)
        self._h = lib.mtldas_train_create(KINDS[kind], self.batch, PRECISIONS[precision])
        if not self._h:
            raise MtldasError("train_create: " + lib.mtldas_last_error().decode())
        self.n_state = lib.mtldas_train_num_state_entries(self._h)
        self.n_params = lib.mtldas_train_num_params(self._h)
        self.workspace_bytes = lib.mtldas_train_workspace_bytes(self._h)
        self.classes = {"MTL": (16, 2), "single_distance": (16,), "single_event": (2,)}[kind]
        self.device, self._ws, self._state, self._arr = None, None, None, None
        if need_device:
            self.device = torch.device("cuda", torch.cuda.current_device()) if device is None else torch.device(device)
            _require_b200(self.device)
            with torch.cuda.device(self.device):
                self._ws = torch.empty(self.workspace_bytes + 256, dtype=torch.uint8, device=self.device)
                ptr = (self._ws.data_ptr() + 255) & ~255
                check(lib.mtldas_train_bind_workspace(self._h, ctypes.c_void_p(ptr), self.workspace_bytes,
                                                      _stream_ptr(self.device)), "train_bind_workspace")

    def __del__(self):
        h = getattr(self, "_h", None)
        if h:
            lib.mtldas_train_destroy(h)
            self._h = None

    def describe(self):
        n = lib.mtldas_train_describe(self._h, None, 0)
        buf = ctypes.create_string_buffer(n)
        lib.mtldas_train_describe(self._h, buf, n)
        return json.loads(buf.value.decode())

    def _state_array(self, state_tensors):
        for t in state_tensors:
            if t.device != self.device or not t.is_contiguous():
                raise MtldasError("state tensors must be contiguous and on the plan's device")
        self._state = list(state_tensors)
        self._arr = (ctypes.c_void_p * self.n_state)(*[t.data_ptr() for t in self._state])

    def forward(self, state_tensors, x):
        """state_tensors: the module's state_dict values in order (parameters fp32, running stats fp32, counters int64).
        Running statistics are updated in place."""
        if not x.is_cuda:
            raise MtldasError("input must be a CUDA tensor; there is no CPU fallback")
        x = x.contiguous().float()
        n = x.shape[0]
        if n > self.batch:
            raise MtldasError(f"batch {n} larger than the plan's {self.batch}")
        self._state_array(state_tensors)
        self._x = x
        outs = [torch.empty((n, c), dtype=torch.float32, device=x.device) for c in self.classes]
        with torch.cuda.device(self.device):
            check(lib.mtldas_train_forward(self._h, self._arr, self.n_state, ctypes.c_void_p(x.data_ptr()), n,
                                           ctypes.c_void_p(outs[0].data_ptr()),
                                           ctypes.c_void_p(outs[1].data_ptr()) if len(outs) > 1 else None,
                                           _stream_ptr(self.device)), "train_forward")
        return tuple(outs)

    def backward(self, douts, flat=None):
        """douts: d loss / d log-probs per task. Returns the flat fp32 gradient (named_parameters order)."""
        douts = [d.contiguous().float() for d in douts]
        if flat is None:
            flat = torch.empty(self.n_params, dtype=torch.float32, device=self.device)
        with torch.cuda.device(self.device):
            check(lib.mtldas_train_backward(self._h, self._arr, self.n_state, ctypes.c_void_p(douts[0].data_ptr()),
                                            ctypes.c_void_p(douts[1].data_ptr()) if len(douts) > 1 else None,
                                            ctypes.c_void_p(flat.data_ptr()), _stream_ptr(self.device)), "train_backward")
        return flat

    @property
    def last_launch_count(self):
        return lib.mtldas_train_last_launch_count(self._h)

    def read_tensor(self, name, nb):
        d = self.describe()
        t = next(t for t in d["tensors"] if t["name"] == name)
        out = torch.empty((nb, t["C"], t["H"], t["W"]), dtype=torch.float32, device=self.device)
        with torch.cuda.device(self.device):
            check(lib.mtldas_train_read_tensor(self._h, name.encode(), nb, ctypes.c_void_p(out.data_ptr()),
                                               _stream_ptr(self.device)), "train_read_tensor")
        return out


def conv2d_nchw(x, weight, bias=None, aux=None, x2=None, stride=1, epilogue=0, precision="bf16",
                aux_phase_split=False):
    """One fused convolution through mtldas_conv2d_nchw (unit-parity entry point)."""
    _require_b200(x.device)
    n, cin, h, w = x.shape
    cout, cin_tot, k, _ = weight.shape
    cin2 = x2.shape[1] if x2 is not None else 0
    assert cin_tot == cin + cin2
    ho, wo = (h, w) if stride == 1 else ((h + 1) // 2, (w + 1) // 2)
    if epilogue == 5:      # fused ReLU + MaxPool2d(2, 2, ceil_mode=True)
        ho, wo = (ho + 1) // 2, (wo + 1) // 2
    out = torch.empty((n, cout, ho, wo), dtype=torch.float32, device=x.device)
    p = lambda t: ctypes.c_void_p(t.contiguous().data_ptr()) if t is not None else None  # noqa: E731
    keep = [t.contiguous().float() if t is not None else None for t in (x, x2, weight, bias, aux)]
    with torch.cuda.device(x.device):
        torch.cuda.synchronize()
        check(lib.mtldas_conv2d_nchw(p(keep[0]), p(keep[1]), p(keep[2]), p(keep[3]), p(keep[4]), p(out), n, cin, cin2,
                                     h, w, cout, k, stride, epilogue, PRECISIONS[precision], int(aux_phase_split)),
              "conv2d_nchw")
    return out


def conv2d_nchw_static(x, weight, bias=None, aux=None, x2=None, stride=1, epilogue=0, precision="bf16",
                       aux_phase_split=False, out=None):
    """The same convolution through mtldas_conv2d_nchw_static: for callers that must not allocate GPU memory
    (streaming loops that cannot take cudaMalloc's device-wide sync, processes whose memory belongs to a pool).
    The library works in a fixed 32 MiB scratch inside its own image; pass `out` to avoid even the result
    allocation.  Asynchronous on the current stream; calls share the scratch, so do not overlap them."""
    _require_b200(x.device)
    n, cin, h, w = x.shape
    cout, cin_tot, k, _ = weight.shape
    cin2 = x2.shape[1] if x2 is not None else 0
    assert cin_tot == cin + cin2
    ho, wo = (h, w) if stride == 1 else ((h + 1) // 2, (w + 1) // 2)
    if out is None:
        out = torch.empty((n, cout, ho, wo), dtype=torch.float32, device=x.device)
    for t in (x, x2, weight, bias, aux, out):
        if t is not None and not (t.is_contiguous() and t.dtype == torch.float32 and t.device == x.device):
            raise MtldasError("conv2d_nchw_static needs contiguous fp32 tensors on one device (it converts nothing)")
    p = lambda t: ctypes.c_void_p(t.data_ptr()) if t is not None else None  # noqa: E731
    with torch.cuda.device(x.device):
        check(lib.mtldas_conv2d_nchw_static(p(x), p(x2), p(weight), p(bias), p(aux), p(out), n, cin, cin2, h, w, cout,
                                            k, stride, epilogue, PRECISIONS[precision], int(aux_phase_split),
                                            _stream_ptr(x.device)), "conv2d_nchw_static")
    return out


def nll_loss(logps, labels, want_grads=True, status=None, out=None):
    """Joint NLLLoss(mean) of the task heads (utils.py:139,361,367) and its gradient through mtldas_nll_loss.

    logps: one or two fp32 [n, classes] log-prob tensors; labels: matching int64 [n] tensors.
    Returns (loss [1] fp32 on the device, [d loss / d logp per task] or None).  Asynchronous on the current stream;
    `status` (int32 [1], optional) is set to 1 by the kernel when a label is out of range; `out` = (loss, grads) from an
    earlier call reuses those buffers (stable addresses keep the training plan's CUDA graph replaying)."""
    dev = logps[0].device
    _require_b200(dev)
    n = logps[0].shape[0]
    for lp, y in zip(logps, labels):
        if not (lp.is_contiguous() and lp.dtype == torch.float32 and y.is_contiguous() and y.dtype == torch.int64
                and lp.shape[0] == n and y.shape[0] == n and y.device == dev and lp.device == dev):
            raise MtldasError("nll_loss needs contiguous fp32 [n, classes] log-probs and int64 [n] labels on one device")
    if out is not None and out[1] is not None and all(g.shape == lp.shape and g.device == dev for g, lp in zip(out[1], logps)):
        loss, grads = out
    else:
        loss = torch.empty(1, dtype=torch.float32, device=dev)
        grads = [torch.empty_like(lp) for lp in logps] if want_grads else None
    p = lambda t: ctypes.c_void_p(t.data_ptr()) if t is not None else None  # noqa: E731
    two = len(logps) > 1
    with torch.cuda.device(dev):
        check(lib.mtldas_nll_loss(p(logps[0]), p(labels[0]), logps[0].shape[1], p(grads[0]) if grads else None,
                                  p(logps[1]) if two else None, p(labels[1]) if two else None,
                                  logps[1].shape[1] if two else 0, p(grads[1]) if (grads and two) else None,
                                  n, p(loss), p(status), _stream_ptr(dev)), "nll_loss")
    return loss, grads


class FlatAdam:
    """torch.optim.Adam(params, lr, weight_decay) of the reference trainer (utils.py:133-134,374) over ONE flat fp32
    parameter buffer through mtldas_adam_step: L2 weight decay folded into the gradient, betas (0.9, 0.999), eps 1e-8,
    bias correction; one launch per step.  `lr` is an attribute (the reference divides it by 1.5 every 5 epochs,
    utils.py:245-247).  State (exp_avg, exp_avg_sq, step) lives here; state_dict()/load_state_dict() keep torch's names."""

    def __init__(self, flat_param, lr=1e-3, betas=(0.9, 0.999), eps=1e-8, weight_decay=0.0):
        if not (flat_param.is_cuda and flat_param.dtype == torch.float32 and flat_param.is_contiguous()):
            raise MtldasError("FlatAdam needs one contiguous fp32 CUDA buffer (see ddp.flatten_parameters_)")
        _require_b200(flat_param.device)
        self.param = flat_param
        self.lr, self.betas, self.eps, self.weight_decay = float(lr), (float(betas[0]), float(betas[1])), float(eps), float(weight_decay)
        self.exp_avg = torch.zeros_like(flat_param.detach())
        self.exp_avg_sq = torch.zeros_like(flat_param.detach())
        self.step_count = 0

    def zero_grad(self, set_to_none=False):
        if self.param.grad is not None and not set_to_none:
            self.param.grad.zero_()
        elif set_to_none:
            self.param.grad = None

    def step(self, grad=None):
        g = self.param.grad if grad is None else grad
        if g is None:
            raise MtldasError("FlatAdam.step: no gradient")
        if not (g.is_cuda and g.dtype == torch.float32 and g.is_contiguous() and g.numel() == self.param.numel()):
            raise MtldasError("FlatAdam.step: the gradient must be a contiguous fp32 CUDA buffer of the parameters' size")
        self.step_count += 1
        dev = self.param.device
        with torch.cuda.device(dev):
            check(lib.mtldas_adam_step(ctypes.c_void_p(self.param.data_ptr()), ctypes.c_void_p(g.data_ptr()),
                                       ctypes.c_void_p(self.exp_avg.data_ptr()), ctypes.c_void_p(self.exp_avg_sq.data_ptr()),
                                       self.param.numel(), self.step_count, self.lr, self.betas[0], self.betas[1], self.eps,
                                       self.weight_decay, _stream_ptr(dev)), "adam_step")

    def state_dict(self):
        return {"step": self.step_count, "exp_avg": self.exp_avg, "exp_avg_sq": self.exp_avg_sq, "lr": self.lr}

    def load_state_dict(self, sd):
        self.step_count = int(sd["step"])
        self.exp_avg.copy_(sd["exp_avg"])
        self.exp_avg_sq.copy_(sd["exp_avg_sq"])
        self.lr = float(sd.get("lr", self.lr))


class EvalMetrics:
    """Validation bookkeeping of utils.py:274-291,313 kept on the device (mtldas_eval_metrics): per-task NLL sums,
    argmax predictions (first maximum wins, like torch.max), correct counts and confusion matrices accumulate over the
    batches of an epoch; result() is the ONE device-to-host read (the reference syncs once per sample)."""

    def __init__(self, classes=(16, 2), device=None):
        self.classes = tuple(int(c) for c in classes)
        self.device = torch.device("cuda", torch.cuda.current_device()) if device is None else torch.device(device)
        _require_b200(self.device)
        self.counts = torch.zeros(4, dtype=torch.int64, device=self.device)
        self.loss_sums = torch.zeros(2, dtype=torch.float64, device=self.device)
        self.confusion = [torch.zeros(c * c, dtype=torch.int64, device=self.device) for c in self.classes]
        self.batches = 0
        self.loss_mean_sums = torch.zeros(2, dtype=torch.float64, device=self.device)

    def reset(self):
        for t in (self.counts, self.loss_sums, self.loss_mean_sums, *self.confusion):
            t.zero_()
        self.batches = 0

    def update(self, logps, labels, want_pred=False):
        """logps: the module's outputs for one batch (fp32 [n, classes_t]); labels: int64 [n] per task."""
        n = logps[0].shape[0]
        for lp, y, c in zip(logps, labels, self.classes):
            if not (lp.is_cuda and lp.dtype == torch.float32 and lp.is_contiguous() and lp.shape == (n, c)
                    and y.dtype == torch.int64 and y.is_contiguous() and y.shape[0] == n and y.device == lp.device):
                raise MtldasError("EvalMetrics.update needs contiguous fp32 [n, classes] log-probs and int64 [n] labels")
        preds = [torch.empty(n, dtype=torch.int64, device=self.device) for _ in logps] if want_pred else [None] * len(logps)
        p = lambda t: ctypes.c_void_p(t.data_ptr()) if t is not None else None  # noqa: E731
        two = len(logps) > 1
        before = self.loss_sums.clone()
        with torch.cuda.device(self.device):
            check(lib.mtldas_eval_metrics(p(logps[0]), p(labels[0]), self.classes[0], p(preds[0]),
                                          p(logps[1]) if two else None, p(labels[1]) if two else None,
                                          self.classes[1] if two else 0, p(preds[1]) if two else None, n,
                                          p(self.counts), p(self.loss_sums), p(self.confusion[0]),
                                          p(self.confusion[1]) if two else None, _stream_ptr(self.device)), "eval_metrics")
        # the reference adds criterion(out, y) = the batch MEAN to testloss (utils.py:274-278): keep that sum as well
        self.loss_mean_sums += (self.loss_sums - before) / max(n, 1)
        self.batches += 1
        return preds if want_pred else None

    def result(self):
        counts = self.counts.cpu().tolist()
        seen = max(counts[2], 1)
        out = {"seen": counts[2], "bad_labels": counts[3],
               "accuracy": [counts[t] / seen for t in range(len(self.classes))],
               "nll_mean": [float(v) / seen for v in self.loss_sums.cpu().tolist()][:len(self.classes)],
               "sum_of_batch_means": self.loss_mean_sums.cpu().tolist()[:len(self.classes)],
               "confusion": [m.view(c, c).cpu() for m, c in zip(self.confusion, self.classes)]}
        return out


def umma_selftest(n_cols=16, row_shift=0, swap=0):
    return float(lib.mtldas_umma_selftest(n_cols, row_shift, swap))
