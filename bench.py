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
    ap.add_argument("--train-batch", type=int, default=256, help="training patches per GPU per step")
    return ap.parse_args()


# ------------------------------------------------------------------------------------------------
# CPU arm: the oracle port on the host cores
# ------------------------------------------------------------------------------------------------
def cpu_samples_per_s(sample, steps, warmup):
    import torch
    from oracle import mtl_oracle as O
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    sd = O.synth_state_dict("MTL", seed=0)
    x, _, _ = O.synth_patches(sample, seed=1234)
    with torch.no_grad():
        for _ in range(warmup):
            O.forward(sd, x, "MTL", training=False)
        t0 = time.perf_counter()
        for _ in range(steps):
            O.forward(sd, x, "MTL", training=False)
        dt = time.perf_counter() - t0
    return sample * steps / dt, dt / steps * 1e3, cores


def cpu_train_samples_per_s(batch, steps, warmup):
    """Reference training step on the host cores: zero_grad, forward, NLL+NLL, backward (utils.py:359-373)."""
    import torch
    from oracle import mtl_oracle as O
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    sd = O.synth_state_dict("MTL", seed=0)
    x, yd, ye = O.synth_patches(batch, seed=1234)
    for _ in range(warmup):
        O.loss_and_grads(sd, x, yd, ye, "MTL")
    t0 = time.perf_counter()
    for _ in range(steps):
        O.loss_and_grads(sd, x, yd, ye, "MTL")
    dt = time.perf_counter() - t0
    return batch * steps / dt, dt / steps * 1e3, cores


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    val, ms, cores = cpu_samples_per_s(args.cpu_sample, args.steps, args.warmup)
    sample = f"{args.cpu_sample} patches/step x {args.steps} steps, eval forward, torch CPU fp32, {cores} threads"
    line = {
        "impl": "reference", "metric": METRIC, "value": val, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "f32", "data": "synthetic",
        "config": {"workload": "MTL (model A) inference, synthetic DAS patches [1,100,250] (bounded CPU sample)",
                   "batch_per_step": args.cpu_sample},
        "cpu_baseline": {"value": val, "unit": UNIT, "cores": cores, "kind": "port", "sample": sample},
        "e2e": {"value": val, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    tval, tms, _ = cpu_train_samples_per_s(32, max(2, min(args.steps, 5)), 1)
    line["train"] = {"value": tval, "unit": UNIT, "ms_per_step": tms, "batch": 32,
                     "what": "reference CPU training step fwd+bwd (BASELINE.json configs[0]), oracle port"}
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
        for g in op["groups"]:
            for s in g["segs"]:
                L = layers[s["layer"]]
                ho, wo = op["H"], op["W"]
                hi, wi = (ho, wo) if L["stride"] == 1 else {17: (33, 83), 9: (17, 42), 5: (9, 21)}[ho]
                e += L["cin"] * hi * wi + L["cout"] * ho * wo
        per_op.append(e)
    return per_op


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
    # streaming: [n_channels, n_time] fp32 per rank, window 100 channels x 250 samples, channel stride 10
    n_ch, n_time, stride = 2140, 250 * 20, 10
    si = StreamingInference(O.synth_state_dict("MTL", seed=0), "MTL", args.precision, min(B, args.micro_batch), stride, dev)
    g = torch.Generator(device="cpu").manual_seed(7 + rank)
    slab = torch.randn(n_ch, n_time, generator=g).to(dev)
    nwin = (n_ch - 100) // stride + 1
    ms = timed(lambda: si.run(slab))
    out["streaming"] = {"value": world * nwin * (n_time // 250) / (ms * 1e-3), "unit": "windows/s", "ms_per_step": ms,
                        "workload": "array [%d channels x %d samples] per GPU -> %d windows x %d blocks, stride %d, "
                                    "window gather + MTL forward, windows sharded across ranks, no collective"
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
    return {"value": world * nb * steps / (float(ms.item()) * 1e-3), "unit": UNIT, "ms_per_step": float(ms.item()) / steps,
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
        peak = 6650.0
        peak_src = "fallback"
        try:
            peak = float(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"])
            peak_src = "measured"
        except Exception:
            pass
        ach = conv_bytes / (conv_ms * 1e-3) / 1e9
        # DRAM read+write bytes per conv launch from the ncu --set full pass over this workload
        # (profiles/r1_conv_ncu_full_batch4096.txt: 15 330 MB over the 30 launches); other workloads: not captured
        traffic = 15330e6 / 30 if (B == 4096 and args.precision == "bf16" and passes == 1) else None
        roof = {"bound": "hbm", "achieved": ach, "peak": peak, "unit": "GB/s", "frac": ach / peak, "traffic": traffic,
                "traffic_source": "profiles/r1_conv_ncu_full_batch4096.txt" if traffic else None,
                "algorithmic_bytes_per_launch": conv_bytes / n_conv,
                "kernel": "conv_umma_persist_kernel (+ conv_umma_kernel on the deep layers)" if args.precision == "bf16" else "conv_fp32_kernel",
                "launches_per_step": n_conv, "avg_launch_ms": conv_ms / n_conv,
                "algorithmic_bytes_per_step": conv_bytes, "peak_source": peak_src,
                "share_of_step": conv_ms / sum(per_op_ms),
                "per_op_ms": {(op.get("name") or op["op"]) + f"#{i}": round(m, 4)
                              for i, (m, op) in enumerate(zip(per_op_ms, desc["program"]))}}

    train = measure_train(args, dev, world, rank, barrier)
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
                    "d2h_bytes_per_step": B * sum(plan.classes) * 4},
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
            val, cms, cores = cpu_samples_per_s(args.cpu_sample, steps, 2)
            line["cpu_baseline"] = {"value": val, "unit": UNIT, "cores": cores, "kind": "port",
                                    "sample": f"{args.cpu_sample} patches/step x {steps} steps, eval forward, "
                                              f"torch CPU fp32 oracle port, {cores} threads"}
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
