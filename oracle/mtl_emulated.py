"""Rounding-emulated training step of MTL_Net / Single_Task_Net.  TEST INFRASTRUCTURE ONLY.

The third oracle tier of SURVEY.md section 8c / Appendix E.3: the reference training step

    /root/reference/utils.py:359-373       out = model(x); loss = NLL + NLL; loss.backward()
    /root/reference/model/modelA_MTL.py:127-174 (forward), modelB_singleTask.py:130-178

restated with a HAND-WRITTEN backward pass (no autograd), so that a rounding function can be applied at exactly
the points where the bf16 CUDA path stores a tensor or feeds a tensor-core operand:

  * convolution operands: activations as stored (bf16), weights rounded to bf16; fp32 (here fp64) accumulation; the
    raw convolution output y is stored rounded;
  * BatchNorm statistics, scale/shift, activations, sigmoid, pooling, softmax, loss: full precision arithmetic on
    the stored (rounded) tensors; every activation tensor is stored rounded;
  * backward: every activation gradient and every dy is stored rounded; a gradient tensor that receives several
    contributions is rounded after EACH accumulation, in the order of the CUDA backward program (task sub-nets of
    level 4..1 first, then resblock8..1; within a block left.3, shortcut, left.0); weight / gamma / beta gradients
    are full-precision reductions of the rounded operands.

With ``rnd=None`` nothing is rounded and the function is an independent derivation of the exact backward pass:
tests/test_oracle_emulated.py checks it against autograd (mtl_oracle.loss_and_grads) in float64, which pins the
hand-written backward to the reference's autograd semantics.  All intermediate tensors are returned under the
names of the CUDA training plan ("y:<conv>", "sb3", "dy:<conv>", "d:<tensor>", tasks stacked along channels) so a
GPU test can localise a deviation to one layer.
"""
from __future__ import annotations

from typing import Callable, Dict, Optional

import torch
import torch.nn.functional as F
from torch.nn.grad import conv2d_input, conv2d_weight

from . import mtl_oracle as O


def bf16_round(t: torch.Tensor) -> torch.Tensor:
    return t.to(torch.bfloat16).to(t.dtype)


class _Step:
    def __init__(self, sd, kind, rnd: Optional[Callable], dtype):
        self.dtype = dtype
        self.sd = {k: (v.detach().to(dtype) if v.is_floating_point() else v) for k, v in sd.items()}
        self.kind = kind
        self.tasks = O.model_tasks(kind)
        self.emulate = rnd is not None
        self.R = rnd if rnd is not None else (lambda t: t)
        # per-channel constants are fp32 variables in the kernels (scale, shift, mean, invstd, A, B)
        self.f32 = (lambda t: t.float().to(dtype)) if self.emulate else (lambda t: t)
        self.T: Dict[str, torch.Tensor] = {}     # named intermediates (plan tensor names)
        self.bn: Dict[str, dict] = {}
        self.cv: Dict[str, tuple] = {}
        self.grads: Dict[str, torch.Tensor] = {}
        self.stats: Dict[str, torch.Tensor] = {}
        self.D: Dict[object, torch.Tensor] = {}  # activation gradients
        self.parts: Dict[str, dict] = {}         # name -> {channel offset: tensor}
        self.forced: Optional[Dict[str, torch.Tensor]] = None
        self.errors: Dict[str, dict] = {}

    def put(self, name, val, c0=0, width=None):
        """Register a stored tensor under its CUDA-plan name (channels [c0, c0+C) of it; `width` = channels of the
        plan tensor when the part is narrower, e.g. the padded level-1 mid tensor of a single-task net).
        Teacher forcing: when `forced` holds that tensor (read back from the GPU), the value computed here FROM FORCED
        INPUTS is compared with the GPU's, the error is recorded, and the GPU's value is what flows on -- so every
        recorded error is local to the op that produced the tensor."""
        if self.forced is not None and name in self.forced:
            ref = self.forced[name][:, c0:c0 + val.shape[1]].to(val.dtype)
            e = self.errors.setdefault(name, dict(num=0.0, den=0.0, amax=0.0, emax=0.0, ndiff=0, count=0))
            d = (val - ref)
            e["num"] += float((d * d).sum()); e["den"] += float((val * val).sum())
            e["amax"] = max(e["amax"], float(val.abs().max())); e["emax"] = max(e["emax"], float(d.abs().max()))
            e["ndiff"] += int((d != 0).sum()); e["count"] += d.numel()
            val = ref
        self.parts.setdefault(name, {})[c0] = val
        if width is not None:
            self.parts[name].setdefault("width", width)
        return val

    def tensors(self):
        out = {}
        for name, parts in self.parts.items():
            width = parts.pop("width", None)
            z = torch.cat([parts[c] for c in sorted(parts)], dim=1)
            if width is not None and z.shape[1] < width:
                z = torch.cat((z, torch.zeros(z.shape[0], width - z.shape[1], *z.shape[2:], dtype=z.dtype)), dim=1)
            out[name] = z
        return out

    # ---- forward pieces -----------------------------------------------------------------------
    def conv(self, prefix, a_in, stride, pad, wgrad_input=None):
        """Raw convolution output as stored (modelA_MTL.py:13,17,23,45,48,74: nn.Conv2d; the pre-BN bias only shifts
        running_mean, see bn())."""
        w = self.R(self.sd[prefix + ".weight"])
        y = self.R(F.conv2d(a_in, w, None, stride=stride, padding=pad))
        self.cv[prefix] = (a_in if wgrad_input is None else wgrad_input, a_in.shape, w, stride, pad)
        return y

    def bn_fwd(self, prefix, conv_prefix, y):
        """nn.BatchNorm2d in train mode on the stored y (modelA_MTL.py:15,18,24,46,49,75,102)."""
        n = y.numel() // y.shape[1]
        mean = y.mean(dim=(0, 2, 3))
        var = (y * y).mean(dim=(0, 2, 3)) - mean * mean
        var = var.clamp_min(0)
        invstd = torch.rsqrt(var + O.BN_EPS)
        g, b = self.sd[prefix + ".weight"], self.sd[prefix + ".bias"]
        scale = self.f32(g * invstd)
        shift = self.f32(b - mean * g * invstd)
        cb = self.sd.get(conv_prefix + ".bias")
        self.stats[prefix + ".running_mean"] = (1 - O.BN_MOMENTUM) * self.sd[prefix + ".running_mean"] + \
            O.BN_MOMENTUM * (mean + (cb if cb is not None else 0))
        self.stats[prefix + ".running_var"] = (1 - O.BN_MOMENTUM) * self.sd[prefix + ".running_var"] + \
            O.BN_MOMENTUM * var * (n / max(n - 1, 1))
        self.stats[prefix + ".num_batches_tracked"] = self.sd[prefix + ".num_batches_tracked"] + 1
        self.bn[prefix] = dict(y=y, mean=self.f32(mean), invstd=self.f32(invstd), scale=scale, shift=shift, n=n)
        return y * scale.view(1, -1, 1, 1) + shift.view(1, -1, 1, 1)

    # ---- backward pieces ----------------------------------------------------------------------
    def bn_bwd(self, prefix, g):
        """BatchNorm backward: dgamma = sum(g*yhat), dbeta = sum(g), dy = scale*(g - mean(g) - yhat*mean(g*yhat)),
        evaluated the way the kernel does (dy = scale*g + A + B*y with fp32 A, B)."""
        s = self.bn[prefix]
        y, mean, invstd, scale, n = s["y"], s["mean"], s["invstd"], s["scale"], s["n"]
        v = lambda t: t.view(1, -1, 1, 1)  # noqa: E731
        yhat = (y - v(mean)) * v(invstd)
        red_g = g.sum(dim=(0, 2, 3))
        red_gy = (g * yhat).sum(dim=(0, 2, 3))
        self.grads[prefix + ".weight"] = self.f32(red_gy)
        self.grads[prefix + ".bias"] = self.f32(red_g)
        cg, cgy = red_g / n, red_gy / n
        A = self.f32(-scale * (cg - cgy * mean * invstd))
        B = self.f32(-scale * cgy * invstd)
        return self.R(v(scale) * g + v(A) + v(B) * y)

    def conv_bwd(self, prefix, dy, need_input=True):
        a_in, in_shape, w, stride, pad = self.cv[prefix]
        self.grads[prefix + ".weight"] = conv2d_weight(a_in, w.shape, dy, stride=stride, padding=pad)
        if (prefix + ".bias") in self.sd:       # a bias in front of train-mode BatchNorm has exactly zero gradient
            self.grads[prefix + ".bias"] = torch.zeros_like(self.sd[prefix + ".bias"])
        return conv2d_input(in_shape, w, dy, stride=stride, padding=pad) if need_input else None

    def acc(self, key, val):
        """First contribution overwrites, later ones accumulate onto the STORED gradient (rounded each time)."""
        self.D[key] = self.R(val) if key not in self.D else self.R(self.D[key] + val)


def train_step(sd, x, yd, ye, kind: str = "MTL", rnd: Optional[Callable] = bf16_round, dtype=torch.float64,
               round_input: bool = True, forced: Optional[Dict[str, torch.Tensor]] = None, return_errors: bool = False):
    """One training-step forward + backward (no optimizer).  Returns (outs, loss, grads, new_stats, tensors)
    [+ per-tensor teacher-forcing errors when return_errors].

    rnd=bf16_round emulates the bf16 tensor-core path (round_input: the tcgen05 stem feeds the patches as bf16;
    its weight gradient still reads the fp32 patches); rnd=None is the exact computation.
    forced: {plan tensor name: [n,C,H,W] tensor read back from the GPU}: see _Step.put."""
    S = _Step(sd, kind, rnd, dtype)
    S.forced = forced
    R, put = S.R, S.put
    tasks = S.tasks
    nt = len(tasks)
    relu = torch.relu
    x = x.to(dtype)

    # ---------------- forward: trunk (modelA_MTL.py:128-139) ----------------
    xs = R(x) if (S.emulate and round_input) else x
    y = put("y:stem", S.conv("conv1.0", xs, 3, 2, wgrad_input=x))
    pre = {"stem": S.bn_fwd("conv1.1", "conv1.0", y)}
    prev = put("stem", R(relu(pre["stem"])))
    sb = []
    for i, (ci, co, s) in enumerate(O.RESBLOCKS, start=1):
        rb = f"resblock{i}"
        down = s != 1 or ci != co
        y1 = put(f"y:{rb}.left.0", S.conv(rb + ".left.0", prev, s, 1))
        pre[rb + ".mid"] = S.bn_fwd(rb + ".left.1", rb + ".left.0", y1)
        mid = put(rb + ".mid", R(relu(pre[rb + ".mid"])))
        if down:
            ysc = put(f"y:{rb}.shortcut.0", S.conv(rb + ".shortcut.0", prev, s, 0))
        y3 = put(f"y:{rb}.left.3", S.conv(rb + ".left.3", mid, 1, 1))
        b3 = S.bn_fwd(rb + ".left.4", rb + ".left.3", y3)
        other = S.bn_fwd(rb + ".shortcut.1", rb + ".shortcut.0", ysc) if down else prev
        pre[f"sb{i-1}"] = b3 + other
        prev = put(f"sb{i-1}", R(relu(pre[f"sb{i-1}"])))
        sb.append(prev)

    # ---------------- forward: task sub-networks (modelA_MTL.py:142-163) ----------------
    a_prev = [None] * nt
    sig = {}
    pool_idx = {}
    amask = [[None] * nt for _ in range(4)]
    for l in range(4):
        an = O.ATT_NAMES[l]
        cm, co = O.ATT_CH[l][1], O.ATT_CH[l][2]
        cmw = 16 * nt if l == 0 else None          # level 1: 8 mid channels per task inside a 16-channel-granular tensor
        for t in range(nt):
            u = sb[0] if l == 0 else torch.cat((sb[2 * l], a_prev[t]), dim=1)
            y0 = put(f"y:{an}.0", S.conv(f"{an}.{t}.0", u, 1, 0), cm * t, cmw)
            pre[(an, t, 0)] = S.bn_fwd(f"{an}.{t}.1", f"{an}.{t}.0", y0)
            mid = put(f"att{l+1}.mid", R(relu(pre[(an, t, 0)])), cm * t, cmw)
            y3 = put(f"y:{an}.3", S.conv(f"{an}.{t}.3", mid, 1, 1), co * t)
            sg = torch.sigmoid(S.bn_fwd(f"{an}.{t}.4", f"{an}.{t}.3", y3))
            sig[(l, t)] = sg
            amask[l][t] = put(f"a{l+1}", R(sg * sb[2 * l + 1]), co * t)
            if l < 3:
                op = f"output_layer{l+1}.{t}"
                oc = O.OUT_CH[l][1]
                yo = put(f"y:output_layer{l+1}", S.conv(op + ".0", amask[l][t], 1, 1), oc * t)
                r = relu(S.bn_fwd(op + ".1", op + ".0", yo))
                pre[(op, "r")] = r
                p, idx = F.max_pool2d(r, kernel_size=2, stride=2, ceil_mode=True, return_indices=True)
                pool_idx[(l, t)] = idx
                a_prev[t] = put(f"p{l+1}", R(p), oc * t)

    # ---------------- heads + loss (modelA_MTL.py:165-172, utils.py:139,361,367) ----------------
    outs = []
    for t, task in enumerate(tasks):
        g = amask[3][t].mean(dim=(2, 3))
        k = O.CH[-1] // O.N_CLASSES[task]
        z = g.reshape(g.shape[0], O.N_CLASSES[task], k).mean(dim=2)
        outs.append(S.f32(F.log_softmax(z, dim=1)))
    labels = O.task_labels(kind, yd, ye)
    n = x.shape[0]
    loss = sum(-(o.gather(1, y_.long().view(-1, 1)).sum() / n) for o, y_ in zip(outs, labels))

    # ---------------- backward: heads ----------------
    D = S.D
    for t, task in enumerate(tasks):
        dlogp = torch.zeros_like(outs[t])
        dlogp.scatter_(1, labels[t].long().view(-1, 1), -1.0 / n)
        k = O.CH[-1] // O.N_CLASSES[task]
        hw = amask[3][t].shape[2] * amask[3][t].shape[3]
        dz = (dlogp - torch.exp(outs[t]) * dlogp.sum(dim=1, keepdim=True)) / (hw * k)
        D[("a", 3, t)] = R(dz.repeat_interleave(k, dim=1)[:, :, None, None].expand_as(amask[3][t]).contiguous())

    # ---------------- backward: task sub-networks, level 4 .. 1 ----------------
    for l in range(3, -1, -1):
        an = O.ATT_NAMES[l]
        cm, co = O.ATT_CH[l][1], O.ATT_CH[l][2]
        cmw = 16 * nt if l == 0 else None
        cs = O.CH[l + 1] if l > 0 else 0                       # channels of the shared half of the concat
        if l < 3:
            oc = O.OUT_CH[l][1]
            for t in range(nt):
                op = f"output_layer{l+1}.{t}"
                r = pre[(op, "r")]
                dz = put(f"d:p{l+1}", D[("p", l, t)], oc * t)
                g = F.max_unpool2d(dz, pool_idx[(l, t)], kernel_size=2, stride=2, output_size=r.shape[-2:]) * (r > 0)
                dy = put(f"dy:output_layer{l+1}", S.bn_bwd(op + ".1", g), oc * t)
                D[("a", l, t)] = R(S.conv_bwd(op + ".0", dy))
        dfeat = 0
        dmids = []
        for t in range(nt):
            dz = put(f"d:a{l+1}", D[("a", l, t)], co * t)
            sg = sig[(l, t)]
            g = dz * sb[2 * l + 1] * sg * (1 - sg)
            dfeat = dfeat + dz * sg
            dy3 = put(f"dy:{an}.3", S.bn_bwd(f"{an}.{t}.4", g), co * t)
            dmids.append(put(f"d:att{l+1}.mid", R(S.conv_bwd(f"{an}.{t}.3", dy3)), cm * t, cmw))
        S.acc(("sb", 2 * l + 1), dfeat)
        shared = 0
        for t in range(nt):
            g = dmids[t] * (pre[(an, t, 0)] > 0)
            dy0 = put(f"dy:{an}.0", S.bn_bwd(f"{an}.{t}.1", g), cm * t, cmw)
            din = S.conv_bwd(f"{an}.{t}.0", dy0)
            if l == 0:
                shared = shared + din
            else:
                shared = shared + din[:, :cs]
                D[("p", l - 1, t)] = R(din[:, cs:])
        S.acc(("sb", 2 * l), shared)

    # ---------------- backward: trunk, resblock8 .. 1, stem ----------------
    for i in range(8, 0, -1):
        rb = f"resblock{i}"
        ci, co, s = O.RESBLOCKS[i - 1]
        down = s != 1 or ci != co
        prev_key = ("sb", i - 2) if i > 1 else "stem"
        dz = put(f"d:sb{i-1}", D[("sb", i - 1)])
        g = dz * (pre[f"sb{i-1}"] > 0)
        if not down:
            S.acc(prev_key, g)                                    # identity shortcut: d residual = g
        dy3 = put(f"dy:{rb}.left.3", S.bn_bwd(rb + ".left.4", g))
        dysc = put(f"dy:{rb}.shortcut.0", S.bn_bwd(rb + ".shortcut.1", g)) if down else None
        dmid = put(f"d:{rb}.mid", R(S.conv_bwd(rb + ".left.3", dy3)))
        if down:
            S.acc(prev_key, S.conv_bwd(rb + ".shortcut.0", dysc))
        g1 = dmid * (pre[rb + ".mid"] > 0)
        dy1 = put(f"dy:{rb}.left.0", S.bn_bwd(rb + ".left.1", g1))
        S.acc(prev_key, S.conv_bwd(rb + ".left.0", dy1))
    dz = put("d:stem", D["stem"])
    g = dz * (pre["stem"] > 0)
    dy = put("dy:stem", S.bn_bwd("conv1.1", g))
    S.conv_bwd("conv1.0", dy, need_input=False)

    grads = {k: S.grads[k] for k in O.param_names(kind)}
    res = (tuple(outs), loss, grads, S.stats, S.tensors())
    if return_errors:
        errs = {k: dict(rel=(e["num"] / max(e["den"], 1e-300)) ** 0.5, max_abs=e["emax"], amax=e["amax"],
                        frac_diff=e["ndiff"] / max(e["count"], 1)) for k, e in S.errors.items()}
        res += (errs,)
    return res
