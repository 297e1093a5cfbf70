#!/usr/bin/env python
"""bench.py -- MTL samples/s on synthetic DAS patches (BASELINE.json metric).

    python bench.py --gpus N --steps K --warmup W [--impl reference] [--batch B] [--precision bf16|fp32]

Workload (config.workload): BASELINE.json configs[1] -- MTL (model A) inference, batch 4096 synthetic
patches per GPU, eval-mode BatchNorm, log-probs for both tasks.  A "step" is one forward pass of the hot
path over the batch.  `value` is timed with inputs resident in HBM (410 MB > L2, so no L2 flush is
needed between iterations); `e2e` goes through the C-ABI host entry point (pinned host buffers, H2D
of the patches and D2H of the log-probs inside the timed region).  N > 1: every rank runs its own batch
(window/batch sharding, no data-path collective) -> weak scaling; time = max over ranks (device events).

`--impl reference` times the reference's CPU path (the oracle port: the same torch-CPU ATen ops the
reference's nn.Module executes; the reference is pure Python so there is nothing to compile into
oracle/_ref) on the host cores, on a bounded sample of the same workload.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
PKG = os.path.join(ROOT, "mtl-das.pytorch_b200")
for p in (ROOT, PKG):
    if p not in sys.path:
        sys.path.insert(0, p)

METRIC = "MTL samples/sec (train fwd+bwd; inference) at 1/2/4/8 B200 vs CPU ref"
UNIT = "samples/s"
PATCH_BYTES = 100 * 250 * 4


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--batch", type=int, default=4096, help="patches per GPU per step")
    ap.add_argument("--micro-batch", type=int, default=int(os.environ.get("MTLDAS_MICRO_BATCH", "4096")))
    ap.add_argument("--precision", default="bf16", choices=["bf16", "fp32"])
    ap.add_argument("--cpu-sample", type=int, default=256, help="patches per CPU-baseline step")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-extras", action="store_true", help="skip the model B / streaming side measurements")
    ap.add_argument("--model-c-batch", type=int, default=2048, help="patches per GPU per step of the model C side measurement")
    ap.add_argument("--no-train", action="store_true", help="skip the training-step measurement (kernel tuning runs)")
    ap.add_argument("--train-batch", type=int, default=256, help="training patches per GPU per step")
    return ap.parse_args()


# ------------------------------------------------------------------------------------------------
# CPU arm: the oracle port on the host cores
# ------------------------------------------------------------------------------------------------
def find_reference():
    """The unmodified reference tree, when this machine has it ($MTLDAS_REF, /root/reference, baseline/_ref).  It exists
    in the build container only (a Python reference cannot travel to the GPU box), so on the box the CPU arm is the
    oracle port: the same ATen CPU ops called from oracle/mtl_oracle.py."""
    for cand in (os.environ.get("MTLDAS_REF"), "/root/reference", os.path.join(ROOT, "baseline", "_ref")):
        if cand and os.path.isfile(os.path.join(cand, "model", "modelA_MTL.py")):
            return cand
    return None


def cpu_model():
    """(kind, eval_forward(x), train_step(x, yd, ye)) on the host cores; kind = "reference" | "port"."""
    import torch
    from oracle import mtl_oracle as O
    sd = O.synth_state_dict("MTL", seed=0)
    ref = find_reference()
    if ref:
        import importlib.util
        spec = importlib.util.spec_from_file_location("_reference_modelA_MTL", os.path.join(ref, "model", "modelA_MTL.py"))
        mod = importlib.util.module_from_spec(spec)
        spec.loader.exec_module(mod)
        net = mod.MTL_Net()
        net.load_state_dict(sd, strict=True)
        crit = torch.nn.NLLLoss()

        def fwd(x):
            net.eval()
            return net(x)

        def train(x, yd, ye):       # utils.py:344,359-373 without the optimizer
            net.train()
            net.zero_grad()
            o1, o2 = net(x)
            (crit(o1, yd.long()) + crit(o2, ye.long())).backward()
        return "reference", fwd, train
    return ("port", lambda x: O.forward(sd, x, "MTL", training=False),
            lambda x, yd, ye: O.loss_and_grads(sd, x, yd, ye, "MTL"))


def cpu_samples_per_s(sample, steps, warmup):
    import torch
    from oracle import mtl_oracle as O
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    kind, fwd, _ = cpu_model()
    x, _, _ = O.synth_patches(sample, seed=1234)
    with torch.no_grad():
        for _ in range(warmup):
            fwd(x)
        t0 = time.perf_counter()
        for _ in range(steps):
            fwd(x)
        dt = time.perf_counter() - t0
    return sample * steps / dt, dt / steps * 1e3, cores, kind


def cpu_train_samples_per_s(batch, steps, warmup):
    """Reference training step on the host cores: zero_grad, forward, NLL+NLL, backward (utils.py:359-373)."""
    import torch
    from oracle import mtl_oracle as O
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    kind, _, train = cpu_model()
    x, yd, ye = O.synth_patches(batch, seed=1234)
    for _ in range(warmup):
        train(x, yd, ye)
    t0 = time.perf_counter()
    for _ in range(steps):
        train(x, yd, ye)
    dt = time.perf_counter() - t0
    return batch * steps / dt, dt / steps * 1e3, cores, kind


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    val, ms, cores, kind = cpu_samples_per_s(args.cpu_sample, args.steps, args.warmup)
    what = "the reference's own MTL_Net (unmodified modules)" if kind == "reference" else "oracle port (same ATen CPU ops)"
    sample = f"{args.cpu_sample} patches/step x {args.steps} steps, eval forward, torch CPU fp32, {cores} threads, {what}"
    line = {
        "impl": "reference", "metric": METRIC, "value": val, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "f32", "data": "synthetic",
        "config": {"workload": "MTL (model A) inference, synthetic DAS patches [1,100,250] (bounded CPU sample)",
                   "batch_per_step": args.cpu_sample},
        "cpu_baseline": {"value": val, "unit": UNIT, "cores": cores, "kind": kind, "sample": sample},
        "e2e": {"value": val, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    tval, tms, _, _ = cpu_train_samples_per_s(32, max(2, min(args.steps, 5)), 1)
    line["train"] = {"value": tval, "unit": UNIT, "ms_per_step": tms, "batch": 32,
                     "what": "reference CPU training step fwd+bwd (BASELINE.json configs[0]), " + what}
    print(json.dumps(line), flush=True)


# ------------------------------------------------------------------------------------------------
# clocks
# ------------------------------------------------------------------------------------------------
class ClockSampler:
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.rows, self.proc = [], None
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--id={index}", f"--query-gpu={self.Q}",
                                          "--format=csv,noheader,nounits", "-lms", "20"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append((time.perf_counter(), line.strip()))

    def stop(self, t0, t1):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        sm, mx, reasons = [], None, set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for t, line in self.rows:
            f = [s.strip() for s in line.split(",")]
            if len(f) < 8:
                continue
            try:
                mhz, mx = float(f[1]), float(f[2])
            except ValueError:
                continue
            if t0 - 0.05 <= t <= t1 + 0.05:
                sm.append(mhz)
                for nme, v in zip(names, f[4:8]):
                    if v == "Active":
                        reasons.add(nme)
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": mx, "reasons": sorted(reasons),
                "samples": len(sm)}


# ------------------------------------------------------------------------------------------------
# algorithmic bytes of the conv kernel launches (each reference conv reads its input once
# and writes its output once, bf16 storage, weights ignored)
# ------------------------------------------------------------------------------------------------
def conv_algorithmic_elems_per_sample(desc):
    layers = {L["conv"]: L for L in desc["layers"]}
    per_op = []
    for op in desc["program"]:
        if op["op"] != "conv":
            per_op.append(0)
            continue
        e = 0
        seen = set()
        for g in op["groups"]:
            for s in g["segs"]:
                if s["layer"] in seen:     # N-split groups of one reference conv: its input / output count once
                    continue
                seen.add(s["layer"])
                L = layers[s["layer"]]
                ho, wo = op["H"], op["W"]
                hi, wi = (ho, wo) if L["stride"] == 1 else {17: (33, 83), 9: (17, 42), 5: (9, 21)}[ho]
                e += L["cin"] * hi * wi + L["cout"] * ho * wo
        per_op.append(e)
    return per_op


def conv_flops_per_sample(desc):
    """2 * MACs per sample of every conv op of the program (for the per-op bound max(FLOP / P, bytes / BW))."""
    layers = {L["conv"]: L for L in desc["layers"]}
    per_op = []
    for op in desc["program"]:
        f, seen = 0.0, set()
        if op["op"] == "conv":
            for g in op["groups"]:
                for s in g["segs"]:
                    if s["layer"] in seen:
                        continue
                    seen.add(s["layer"])
                    L = layers[s["layer"]]
                    f += 2.0 * L["cin"] * L["cout"] * L["k"] * L["k"] * op["H"] * op["W"]
        per_op.append(f)
    return per_op


def hbm_peak():
    try:
        return float(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"]), "measured"
    except Exception:
        return 6650.0, "fallback"


def profiled_traffic(version, batch, precision, passes):
    """DRAM read+write bytes per conv launch from the committed `ncu --set full` summary (profiles/r2_traffic.json,
    written by tools/ncu_traffic_summary.py).  Only trusted when it was captured on THIS build of the kernels and this
    workload; otherwise null (the bench never prints a stale number)."""
    try:
        t = json.load(open(os.path.join(ROOT, "profiles", "r2_traffic.json")))
    except Exception:
        return None, None
    if t.get("mtldas_version") != version or t.get("batch") != batch or t.get("precision") != precision or passes != 1:
        return None, "profiles/r2_traffic.json is from build %r, batch %r: not this run" % (t.get("mtldas_version"), t.get("batch"))
    return float(t["conv_dram_bytes_per_step"]) / float(t["conv_launches_per_step"]), t.get("source")


def per_op_fracs(desc, per_op_ms, elems, batch, precision, hbm_gbs):
    """measured / bound per conv op, bound = max(FLOP / sustained bf16 peak, algorithmic bytes / measured HBM peak)."""
    try:
        tf = float(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["bf16_tflops_sustained"])
    except Exception:
        tf = 1400.0
    flops = conv_flops_per_sample(desc)
    out = {}
    for i, (ms, op) in enumerate(zip(per_op_ms, desc["program"])):
        if op["op"] != "conv" or ms <= 0:
            continue
        t_f = flops[i] * batch / (tf * 1e12) * 1e3
        t_b = elems[i] * batch * (2 if precision == "bf16" else 4) / (hbm_gbs * 1e9) * 1e3
        out[op["name"] + f"#{i}"] = round(max(t_f, t_b) / ms, 3)
    return out


def measure_extras(args, dev, world, rank, barrier):
    """BASELINE.json configs[3] and configs[4], reported next to the headline (same timing rules, fewer steps):
    model B (single_distance / single_event) inference on the same kernels, and streaming inference over a synthetic
    DAS array sharded by fibre-channel windows (no collective)."""
    import torch
    import torch.distributed as dist
    from mtldas_b200 import EvalPlan, synth as O
    from mtldas_b200.stream_infer import StreamingInference
    steps = max(3, min(args.steps, 10))

    def timed(fn):
        for _ in range(3):
            fn()
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(steps):
            fn()
        e1.record()
        barrier()
        ms = torch.tensor([e0.elapsed_time(e1)], device=dev)
        if world > 1:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        return float(ms.item()) / steps

    out = {}
    B = args.batch
    base, _, _ = O.synth_patches(64, seed=99 + rank)
    x = base.repeat((B + 63) // 64, 1, 1, 1)[:B].contiguous().to(dev)
    for kind in ("single_distance", "single_event"):
        plan = EvalPlan(kind, min(B, args.micro_batch), args.precision, dev)
        plan.pack({k: v.to(dev) for k, v in O.synth_state_dict(kind, seed=0).items()})
        outs = [torch.empty(B, c, device=dev) for c in plan.classes]
        ms = timed(lambda: plan.forward(x, outs))
        out["model_b_" + kind] = {"value": world * B / (ms * 1e-3), "unit": UNIT, "ms_per_step": ms, "batch_per_gpu": B,
                                  "workload": "Single_Task_Net inference, inputs resident"}
        del plan, outs
    # model C (Multi_Classifier = Inception-V3, BASELINE.json configs[4]): eval forward, compute bound -> tensor roofline
    try:
        from mtldas_b200.nets_c import Multi_Classifier
        nc = min(B, args.model_c_batch)
        mc = Multi_Classifier(init_weights=False)
        mc.load_state_dict(O.synth_state_dict_c(seed=0), strict=True)
        mc = mc.to(dev).eval()
        xc = x[:nc].contiguous()
        ms = timed(lambda: mc(xc))
        flop = 2 * 1.109e9 * nc                       # 1.109 GMAC per patch (94 convolutions)
        try:
            pk = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
            peak_tf, src = float(pk["bf16_tflops"]), "measured (burst)"
        except Exception:
            peak_tf, src = 1650.0, "fallback"
        ach = flop / (ms * 1e-3) / 1e12
        out["model_c"] = {"value": world * nc / (ms * 1e-3), "unit": UNIT, "ms_per_step": ms, "batch_per_gpu": nc,
                          "workload": "Multi_Classifier (Inception-V3, 1-channel stem, 32 classes) eval forward, inputs resident",
                          "gpu_launches_per_step": mc.last_launch_count,
                          "roofline": {"bound": "tensor", "achieved": ach, "peak": peak_tf, "unit": "TFLOP/s",
                                       "frac": ach / peak_tf, "traffic": None, "peak_source": src,
                                       "algorithmic_flop_per_step": flop}}
        del mc
    except Exception as e:      # the headline must not die on a side measurement
        out["model_c"] = {"error": repr(e)[:300]}
    # streaming: [n_channels, n_time] fp32 per rank, window 100 channels x 250 samples, channel stride 10
    n_ch, n_time, stride = 2140, 250 * 20, 10
    si = StreamingInference(O.synth_state_dict("MTL", seed=0), "MTL", args.precision, min(B, args.micro_batch), stride, dev)
    g = torch.Generator(device="cpu").manual_seed(7 + rank)
    slab = torch.randn(n_ch, n_time, generator=g).to(dev)
    nwin = (n_ch - 100) // stride + 1
    ms = timed(lambda: si.run(slab))
    out["streaming"] = {"value": world * nwin * (n_time // 250) / (ms * 1e-3), "unit": "windows/s", "ms_per_step": ms,
                        "workload": "array [%d channels x %d samples] per GPU -> %d windows x %d blocks, stride %d, "
                                    "windows read in place by the stem (no patch tensor) + MTL forward, windows sharded across ranks, no collective"
                                    % (n_ch, n_time, nwin, n_time // 250, stride)}
    return out


def measure_train(args, dev, world, rank, barrier):
    """BASELINE.json configs[2]: MTL training, per-GPU batch 256, data parallel with ONE NCCL all-reduce of the flat
    fp32 gradient per step.  A step = weight packing + forward (batch-stat BN) + backward + all-reduce + Adam step,
    through the drop-in module / DataParallelTrainer."""
    import torch
    import torch.distributed as dist
    from mtldas_b200 import MTL_Net, synth as O
    from mtldas_b200.ddp import DataParallelTrainer
    nb = args.train_batch
    m = MTL_Net()
    m.precision = args.precision
    m.load_state_dict(O.synth_state_dict("MTL", seed=0), strict=True)
    m = m.to(dev).train()
    # Adam(lr 1e-3 / 1.5, L2 weight decay 1e-5) as utils.py:133-134,247, stepped over the flattened parameter buffer
    trainer = DataParallelTrainer(m, flat_adam=dict(lr=1e-3 / 1.5, weight_decay=1e-5))
    base, yd, ye = O.synth_patches(64, seed=4321 + rank)
    reps = (nb + 63) // 64
    xh = base.repeat(reps, 1, 1, 1)[:nb].contiguous().pin_memory()
    ydd, yed = yd.repeat(reps)[:nb].to(dev), ye.repeat(reps)[:nb].to(dev)
    x = xh.to(dev)
    for _ in range(3):
        trainer.step(x, [ydd, yed])
    barrier()
    steps = max(3, min(args.steps, 20))
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(steps):
        loss = trainer.step(x, [ydd, yed])
    e1.record()
    barrier()
    ms = torch.tensor([e0.elapsed_time(e1)], device=dev)
    if world > 1:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    # end to end: H2D of the batch from pinned memory + D2H of the loss every step
    # (double-buffered: the copy of step i+1 is issued on a side stream before the loss of step i is read back)
    copy_stream = torch.cuda.Stream(device=dev)
    slots = [torch.empty_like(x), torch.empty_like(x)]
    ready = [torch.cuda.Event(), torch.cuda.Event()]
    consumed = [torch.cuda.Event(), torch.cuda.Event()]

    def upload(i):
        with torch.cuda.stream(copy_stream):
            copy_stream.wait_event(consumed[i & 1])
            slots[i & 1].copy_(xh, non_blocking=True)
            ready[i & 1].record(copy_stream)

    for ev in consumed:
        ev.record()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    upload(0)
    for i in range(steps):
        torch.cuda.current_stream().wait_event(ready[i & 1])
        loss = trainer.step(slots[i & 1], [ydd, yed])
        consumed[i & 1].record()
        if i + 1 < steps:
            upload(i + 1)
        lv = float(loss)
    barrier()
    e2e = torch.tensor([time.perf_counter() - t0], device=dev)
    if world > 1:
        dist.all_reduce(e2e, op=dist.ReduceOp.MAX)
    plan = m._train_plan(dev, nb)
    # every rank must hold the same parameters after the timed steps (one all-reduced gradient, one optimizer rule)
    from mtldas_b200.ddp import replica_checksum
    spread, sums = replica_checksum(m)
    if world > 1 and spread != 0.0:
        raise SystemExit("bench: replicas diverged after %d training steps (checksum spread %g)" % (steps, spread))
    step_s = float(ms.item()) * 1e-3 / steps
    peak, peak_src = hbm_peak()
    # BASELINE.md section 2 / SURVEY 8d, layer-fused byte model of one training step in bf16: 15.91 MB and
    # 1299.56 MFLOP per sample -> bound 393 k samples/s per GPU at the measured peaks (HBM-bound at layer granularity)
    tbytes = 15.91e6 * (2.0 if args.precision == "fp32" else 1.0) * nb
    train_roof = {"bound": "hbm", "achieved": tbytes / step_s / 1e9, "peak": peak, "unit": "GB/s",
                  "frac": tbytes / step_s / 1e9 / peak, "traffic": None, "peak_source": peak_src,
                  "algorithmic_bytes_per_step": tbytes, "algorithmic_flop_per_step": 1299.56e6 * nb,
                  "what": "whole training step (all launches) against the layer-fused byte model; at 256 patches per GPU "
                          "every tensor is L2-resident, so this fraction measures launch-chain latency, not DRAM"}
    return {"value": world * nb * steps / (float(ms.item()) * 1e-3), "unit": UNIT, "ms_per_step": float(ms.item()) / steps,
            "roofline": train_roof, "replica_checksum_spread": spread, "replica_checksum": sums,
            "optimizer": "mtldas_adam_step (one launch over the flat parameter buffer)" if trainer._fused is None else "inside the fused collective kernel",
            "collective": trainer.collective, "fused_collective_unavailable": getattr(trainer, "fused_unavailable", None),
            "steps": steps, "batch_per_gpu": nb, "scaling": "weak",
            "step": "pack weights + forward (batch-stat BN) + backward + all-reduce + Adam",
            "allreduce_bytes_per_step": plan.n_params * 4 if world > 1 else 0,
            "e2e": {"value": world * nb * steps / float(e2e.item()), "unit": UNIT,
                    "h2d_bytes_per_step": nb * PATCH_BYTES, "d2h_bytes_per_step": 4},
            "gpu_launches_per_step": plan.last_launch_count, "last_loss": lv}


def bind_to_gpu_numa_node(dev):
    """Multi-GPU runs: pin this rank's host threads to the CPUs NVML lists as local to its GPU, BEFORE the pinned
    staging buffers are allocated (first touch places them on that NUMA node) -- eight ranks copying 410 MB per step
    from the wrong socket halve the host-to-device rate.  Best effort: returns the CPU count bound, or None."""
    try:
        import pynvml
        import torch
        pynvml.nvmlInit()
        uuid = str(torch.cuda.get_device_properties(dev).uuid)
        if not uuid.startswith("GPU-"):
            uuid = "GPU-" + uuid
        try:
            h = pynvml.nvmlDeviceGetHandleByUUID(uuid.encode())
        except Exception:
            h = pynvml.nvmlDeviceGetHandleByUUID(uuid)
        words = (os.cpu_count() + 63) // 64
        mask = pynvml.nvmlDeviceGetCpuAffinity(h, words)
        cpus = {64 * i + b for i, w in enumerate(mask) for b in range(64) if (w >> b) & 1}
        cpus &= set(os.sched_getaffinity(0))
        if cpus:
            os.sched_setaffinity(0, cpus)
            return len(cpus)
    except Exception:
        pass
    return None


def run_ours(args):
    import torch
    import torch.distributed as dist
    from mtldas_b200 import EvalPlan, synth as O   # product-side synthetic weights / patches generator

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    numa = bind_to_gpu_numa_node(dev) if world > 1 else None
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)

    B = args.batch
    plan = EvalPlan("MTL", args.micro_batch, args.precision, dev)
    sd = O.synth_state_dict("MTL", seed=0)
    plan.pack({k: v.to(dev) for k, v in sd.items()})
    # synthetic patches: 64 distinct oracle patches tiled to the batch (+ per-rank roll so ranks differ)
    base, _, _ = O.synth_patches(64, seed=1234 + rank)
    x_host = base.repeat((B + 63) // 64, 1, 1, 1)[:B].contiguous().pin_memory()
    x = x_host.to(dev)
    outs_host = [torch.empty((B, c), dtype=torch.float32).pin_memory() for c in plan.classes]

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # ---- device-resident timing -------------------------------------------------------------------
    outs = [torch.empty((B, c), dtype=torch.float32, device=dev) for c in plan.classes]
    for _ in range(max(args.warmup, 3)):
        plan.forward(x, outs)
    barrier()
    sampler = ClockSampler(local) if rank == 0 else None
    t_wall0 = time.perf_counter()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    t_host0 = time.perf_counter()
    for _ in range(args.steps):
        plan.forward(x, outs)
    host_ms = (time.perf_counter() - t_host0) * 1e3 / args.steps
    e1.record()
    barrier()
    t_wall1 = time.perf_counter()
    ms = torch.tensor([e0.elapsed_time(e1)], device=dev)
    if world > 1:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    ms_total = float(ms.item())
    launches = plan.last_launch_count * args.steps
    clocks = sampler.stop(t_wall0, t_wall1) if sampler else None

    # ---- end to end through the host entry point ----------------------------------------------------
    for _ in range(2):
        plan.forward_host(x_host, outs_host)
    barrier()
    e2e_steps = max(3, min(args.steps, 10))
    t0 = time.perf_counter()
    for _ in range(e2e_steps):
        plan.forward_host(x_host, outs_host)
    barrier()
    e2e_s = torch.tensor([time.perf_counter() - t0], device=dev)
    if world > 1:
        dist.all_reduce(e2e_s, op=dist.ReduceOp.MAX)
    e2e_val = world * B * e2e_steps / float(e2e_s.item())
    # the same with the patches already bf16 on the host (half the PCIe bytes; bit-identical log-probs)
    e2e_bf16 = None
    if args.precision == "bf16":
        xh_host = x_host.to(torch.bfloat16).pin_memory()
        for _ in range(2):
            plan.forward_host(xh_host, outs_host)
        barrier()
        t0 = time.perf_counter()
        for _ in range(e2e_steps):
            plan.forward_host(xh_host, outs_host)
        barrier()
        tb = torch.tensor([time.perf_counter() - t0], device=dev)
        if world > 1:
            dist.all_reduce(tb, op=dist.ReduceOp.MAX)
        e2e_bf16 = {"value": world * B * e2e_steps / float(tb.item()), "unit": UNIT, "h2d_bytes_per_step": B * PATCH_BYTES // 2,
                    "d2h_bytes_per_step": B * sum(plan.classes) * 4,
                    "what": "mtldas_plan_forward_eval_host_bf16: host patches stored as bf16"}
        del xh_host
    # host-side bound: pinned host -> device copy rate of this rank while all ranks copy (max over ranks of the time)
    barrier()
    t0 = time.perf_counter()
    for _ in range(3):
        x.copy_(x_host, non_blocking=True)
    barrier()
    tc = torch.tensor([time.perf_counter() - t0], device=dev)
    if world > 1:
        dist.all_reduce(tc, op=dist.ReduceOp.MAX)
    h2d_gbs = 3 * B * PATCH_BYTES / float(tc.item()) / 1e9
    e2e_bound = {"h2d_gbs_per_rank": h2d_gbs, "samples_per_s_fp32_upload": world * h2d_gbs * 1e9 / PATCH_BYTES,
                 "samples_per_s_bf16_upload": world * h2d_gbs * 1e9 / (PATCH_BYTES / 2),
                 "what": "pinned host-to-device copy rate with all %d ranks copying at once: the ceiling of e2e" % world}

    # ---- live per-kernel durations for the roofline (CUDA events around every launch) --------------
    roof = None
    if rank == 0:
        desc = plan.describe()
        per_op_ms = plan.profile(x)
        elems = conv_algorithmic_elems_per_sample(desc)
        conv_ms = sum(m for m, op in zip(per_op_ms, desc["program"]) if op["op"] == "conv")
        conv_bytes = sum(elems) * B * (2 if args.precision == "bf16" else 4)
        passes = (B + args.micro_batch - 1) // args.micro_batch
        n_conv = sum(1 for op in desc["program"] if op["op"] == "conv") * passes
        peak, peak_src = hbm_peak()
        ach = conv_bytes / (conv_ms * 1e-3) / 1e9
        from mtldas_b200 import lib
        traffic, traffic_src = profiled_traffic(lib.mtldas_version().decode(), B, args.precision, passes)
        roof = {"bound": "hbm", "achieved": ach, "peak": peak, "unit": "GB/s", "frac": ach / peak, "traffic": traffic,
                "traffic_source": traffic_src,
                "algorithmic_bytes_per_launch": conv_bytes / n_conv,
                "kernel": "conv_umma_persist_kernel (tcgen05 implicit GEMM; + conv_umma_kernel where weights stream)" if args.precision == "bf16" else "conv_fp32_kernel",
                "launches_per_step": n_conv, "avg_launch_ms": conv_ms / n_conv,
                "algorithmic_bytes_per_step": conv_bytes, "peak_source": peak_src,
                "share_of_step": conv_ms / sum(per_op_ms),
                "per_op_frac_of_own_bound": per_op_fracs(desc, per_op_ms, elems, B, args.precision, peak),
                "per_op_ms": {(op.get("name") or op["op"]) + f"#{i}": round(m, 4)
                              for i, (m, op) in enumerate(zip(per_op_ms, desc["program"]))}}

    train = measure_train(args, dev, world, rank, barrier) if not args.no_train else None
    extras = measure_extras(args, dev, world, rank, barrier) if not args.no_extras else None

    if rank == 0:
        value = world * B * args.steps / (ms_total * 1e-3)
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": max(args.warmup, 3), "ms_per_step": ms_total / args.steps, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None,
            "dtype": "bf16" if args.precision == "bf16" else "f32", "data": "synthetic",
            "config": {"workload": "MTL (model A) inference, eval-mode BN, batch %d synthetic DAS patches [1,100,250] "
                                   "per GPU (BASELINE.json configs[1])" % B,
                       "batch_per_gpu": B, "micro_batch": args.micro_batch, "precision": args.precision,
                       "l2": "inputs (%.0f MB/step) larger than L2; no flush" % (B * PATCH_BYTES / 1e6),
                       "parallelism": "batch-sharded x%d, no collective" % world,
                       "host_cpus_bound_per_rank": numa},
            "e2e": {"value": e2e_val, "unit": UNIT, "h2d_bytes_per_step": B * PATCH_BYTES,
                    "d2h_bytes_per_step": B * sum(plan.classes) * 4,
                    "what": "mtldas_plan_forward_eval_host: fp32 host patches (what the reference's DataLoader holds)"},
            "e2e_bf16_upload": e2e_bf16,
            "e2e_bound": e2e_bound,
            "gpu_launches": launches,
            "host_enqueue_ms_per_step": host_ms,
            "cuda_graph": os.environ.get("MTLDAS_NO_GRAPH", "0") != "1",
            "clocks": clocks,
            "roofline": roof,
            "train": train,
            "extras": extras,
        }
        if world == 1 and not args.no_cpu_baseline:
            steps = 16
            val, cms, cores, kind = cpu_samples_per_s(args.cpu_sample, steps, 2)
            line["cpu_baseline"] = {"value": val, "unit": UNIT, "cores": cores, "kind": kind,
                                    "sample": f"{args.cpu_sample} patches/step x {steps} steps, eval forward, torch CPU "
                                              f"fp32 ({'reference modules' if kind == 'reference' else 'oracle port'}), "
                                              f"{cores} threads"}
            tval, tms, _, _ = cpu_train_samples_per_s(32, 4, 1)
            if line["train"] is not None:
                line["train"]["cpu_baseline"] = {
                    "value": tval, "unit": UNIT, "cores": cores, "kind": kind,
                    "sample": "32 patches/step x 4 steps, train-mode forward + NLL+NLL + backward "
                              "(BASELINE.json configs[0]), torch CPU fp32, %d threads" % cores}
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


def main():
    args = parse()
    if args.gpus > 1 and "RANK" not in os.environ:
        cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={args.gpus}",
               "--master-addr", "127.0.0.1", "--master-port", "29517", os.path.abspath(__file__)] + sys.argv[1:]
        sys.exit(subprocess.call(cmd))
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
