#!/usr/bin/env python
"""bench.py — VGG16 deconvnet projections/s on B200 (BASELINE.json metric), one JSON line.

A "step" is one pass of the hot path over one batch of synthetic input.  Workload at every N (weak scaling, no
data-path collective): each rank takes its own 224x224 synthetic image, runs the forward tape to block4_conv3 once,
projects ALL 512 feature maps of block4_conv3 to pixel space (BASELINE.json configs[1]) and de-processes every
projection to uint8 (reference deepdream.py:483-498, what the reference's only consumer does with a projection,
main.py:72).  `value` is timed with inputs and outputs resident in HBM; `e2e` is the same step through the C-ABI with
HOST buffers (pinned image in, 512 uint8 projections out).  The same step with the raw float32 projections as the
result (`fp32_output`) and in the reference-precision mode (`fp32_mode`) is reported beside it.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--precision bf16|fp32]

N > 1: launched by torchrun, one rank per GPU; weights are broadcast once over NCCL, then no collective touches the
data path.  --impl reference times the CPU port of the reference (the oracle; TensorFlow 1.13 is not installable here)
on the box's host cores, rank 0 only: ONE CPU process against N GPUs.
"""
import argparse
import ctypes
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "vgg16_deconv_projections_per_sec_224"
UNIT = "projections/s"
LAYER = "block4_conv3"
NFEAT = 512
HW = 224
ITEM = HW * HW * 3


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--precision", default=os.environ.get("DCV_BENCH_PRECISION", "bf16"), choices=["bf16", "fp32"])
    ap.add_argument("--layer", default=LAYER)
    ap.add_argument("--nfeat", type=int, default=NFEAT)
    ap.add_argument("--cpu-seconds", type=float, default=12.0, help="CPU work per reference step / cpu_baseline sample")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-dream", action="store_true", help="skip the secondary DeepDream measurement")
    ap.add_argument("--no-extras", action="store_true", help="skip fp32_output / fp32_mode / config4 / config5 / parity")
    ap.add_argument("--dream-size", type=int, default=512)
    return ap.parse_args()


def workload_config(layer, nfeat):
    """The `config` object of BOTH arms (the driver compares them)."""
    return {"workload": "VGG16 224x224 random-init (Glorot, seed 1234), forward to %s + projection of all %d feature maps to "
                        "pixel space + de-process of every projection to uint8, batch 1 per GPU (BASELINE configs[1]); one "
                        "synthetic image per rank" % (layer, nfeat),
            "l2": "no flush needed: each step streams > 6 GB of intermediates, far above the 126 MB L2",
            "parallelism": "replicated weights, image-sharded, no per-step collective"}


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return dict(hbm=d["hbm_gbs"], tf_burst=d["bf16_tflops"], tf_sus=d.get("bf16_tflops_sustained", d["bf16_tflops"]),
                    source="measured")
    return dict(hbm=6650.0, tf_burst=1590.0, tf_sus=1400.0, source="fallback")


class ClockSampler(object):
    """nvidia-smi sampler running during the timed region (B200_PROFILING.md clocks line)."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.proc, self.lines = index, None, []

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q,
                                          "--format=csv,noheader,nounits", "-lms", "20"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._pump, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _pump(self):
        for line in self.proc.stdout:
            self.lines.append(line)

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], 0.0, set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for line in self.lines:
            f = [x.strip() for x in line.split(",")]
            if len(f) < 7:
                continue
            try:
                sm.append(float(f[0])); mx = max(mx, float(f[1]))
            except ValueError:
                continue
            for n, v in zip(names, f[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        sm.sort()
        # median over the busiest half of the samples (idle samples before/after the region are low)
        load = sm[len(sm) // 2:] if sm else []
        return {"sm_mhz": (load[len(load) // 2] if load else None), "sm_max_mhz": mx or None,
                "reasons": sorted(reasons), "samples": len(sm)}


# --------------------------------------------------------------------------------------------- CPU arms
_CPU_TAPE = {}


def cpu_projection_rate(layer, nproj, seed=0, batch=16, deprocess=True):
    """Times the oracle port (vectorised torch-CPU restatement of reference deepdream.py) on the host: forward tape +
    `nproj` projections of `layer` (+ deprocess_image of each), all host threads.  -> (proj/s, seconds, cores)."""
    import torch
    from oracle import keras_shim, restatement as rs
    from deconv_api_b200 import synth
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    if "model" not in _CPU_TAPE:
        _CPU_TAPE["model"] = keras_shim.VGG16(include_top=False, input_shape=(HW, HW, 3))
    model = _CPU_TAPE["model"]
    img = synth.image_batch(1, HW, HW, seed)
    t0 = time.perf_counter()
    tape = rs.FastTape(model, layer, img)
    feats = [f % 512 for f in range(nproj)]
    for done in range(0, nproj, batch):
        proj = tape.project(feats[done:done + batch])
        if deprocess:
            for k in range(proj.shape[0]):
                rs.deprocess_image(proj[k])
    dt = time.perf_counter() - t0
    return nproj / dt, dt, cores


def cpu_sample_size(layer, nfeat, seconds):
    """How many of the `nfeat` projections fit in ~`seconds` of CPU work (the full set whenever it does)."""
    cpu_projection_rate(layer, 8)                                  # warm-up (thread pool, oneDNN primitives)
    rate, _, _ = cpu_projection_rate(layer, 32)
    n = int(rate * seconds) // 32 * 32
    return max(32, min(nfeat, n)) if n < nfeat else nfeat


def reference_faithful_record():
    """The UNMODIFIED reference (app/deepdream.py on the Keras shim, Python pool/unpool loops included) can only run
    where /root/reference exists, i.e. in the build container: profiles/reference_faithful.py measures it there and
    commits the result; this just reports that record."""
    p = os.path.join(ROOT, "profiles", "r2_reference_faithful.json")
    if os.path.exists(p):
        return json.load(open(p))
    return None


def run_reference(args, rank, world):
    if rank != 0:
        return
    sample = cpu_sample_size(args.layer, args.nfeat, args.cpu_seconds)
    rates, secs = [], []
    for _ in range(max(0, min(args.warmup, 1))):
        cpu_projection_rate(args.layer, 32)
    for _ in range(args.steps):
        r, s, cores = cpu_projection_rate(args.layer, sample)
        rates.append(r); secs.append(s)
    value = sample * len(secs) / sum(secs)
    what = ("forward to %s + %d of the %d projections + deprocess_image of each per step, vectorised torch-CPU restatement of "
            "reference deepdream.py on %d threads (TensorFlow 1.13 CPU path unavailable on this box)"
            % (args.layer, sample, args.nfeat, cores))
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": 1e3 * sum(secs) / len(secs) * (args.nfeat / float(sample)),
        "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": workload_config(args.layer, args.nfeat),
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": cores, "kind": "port", "sample": what},
        "cpu_baseline_reference": reference_faithful_record(),
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
        "note": "one CPU process on rank 0 at every --gpus N (the other ranks exit); a ratio against N GPUs is N GPUs vs 1 host",
    }
    print(json.dumps(line), flush=True)


# --------------------------------------------------------------------------------------------- GPU arm
def bind_near_gpu(local_rank):
    """Pin this process (and so the first touch of its pinned buffers) to the CPUs of the GPU's NUMA node."""
    try:
        import torch
        p = torch.cuda.get_device_properties(local_rank)
        path = "/sys/bus/pci/devices/%04x:%02x:%02x.0/" % (p.pci_domain_id, p.pci_bus_id, p.pci_device_id)
        node = int(open(path + "numa_node").read().strip())
        cpus = open(path + "local_cpulist").read().strip()
        if node < 0 or not cpus:
            return {"numa_node": node, "bound": False}
        ids = set()
        for part in cpus.split(","):
            a, _, b = part.partition("-")
            ids.update(range(int(a), int(b or a) + 1))
        ids &= os.sched_getaffinity(0)
        if ids:
            os.sched_setaffinity(0, ids)
        return {"numa_node": node, "bound": bool(ids), "cpus": len(ids)}
    except Exception as e:                                           # VMs often expose no topology
        return {"numa_node": None, "bound": False, "why": type(e).__name__}


def run_ours(args, rank, world, local_rank):
    import torch
    from deconv_api_b200 import synth, vgg16
    torch.cuda.set_device(local_rank)
    numa = bind_near_gpu(local_rank)
    dist = None
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))

    # ---- weights: rank 0 draws them, one NCCL broadcast over NVLink, every rank loads its engine ----
    from deconv_api_b200 import parallel
    model = vgg16.VGG16(weights=None, include_top=False, input_shape=(HW, HW, 3))
    if world > 1:
        parallel.broadcast_model_weights(model, dist, src=0, device=torch.device("cuda", local_rank))
    eng = model.bound_engine(local_rank, upto=args.layer)
    li = eng.index[args.layer]
    feats = np.arange(args.nfeat, dtype=np.int32)
    prec = args.precision
    prec_id = 1 if prec == "bf16" else 0
    U8, F32 = 2, 0

    img = synth.image_batch(1, HW, HW, seed0=rank)                    # this rank's unit of work
    out_elems = args.nfeat * ITEM
    img_dev = torch.from_numpy(img).cuda()
    out_dev = torch.empty(out_elems, dtype=torch.float32, device="cuda")
    out8_dev = torch.empty(out_elems, dtype=torch.uint8, device="cuda")
    img_pin = torch.from_numpy(img).pin_memory()
    out_pin = torch.empty(out_elems, dtype=torch.float32).pin_memory()
    out8_pin = torch.empty(out_elems, dtype=torch.uint8).pin_memory()
    stream = torch.cuda.ExternalStream(eng.stream, device=local_rank)
    fptr = feats.ctypes.data_as(ctypes.c_void_p)

    def step_dev(fmt=U8, p=prec_id):
        eng.forward_dev(img_dev.data_ptr(), 1, HW, HW, li)
        dst = out8_dev if fmt == U8 else out_dev
        eng._ck(eng.lib.dcv_project_fmt_dev(eng.ctx, li, fptr, args.nfeat, 0, p, fmt, ctypes.c_void_p(dst.data_ptr())))

    def step_host(fmt=U8, p=prec_id):
        eng._ck(eng.lib.dcv_forward(eng.ctx, ctypes.c_void_p(img_pin.data_ptr()), 1, HW, HW, li))
        dst = out8_pin if fmt == U8 else out_pin
        eng._ck(eng.lib.dcv_project_fmt(eng.ctx, li, fptr, args.nfeat, 0, p, fmt, ctypes.c_void_p(dst.data_ptr())))

    def barrier():
        torch.cuda.synchronize()
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()

    def reduce_max(x):
        if dist is None:
            return x
        t = torch.tensor([x], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    def time_dev(fn, steps, warm):
        for _ in range(warm):
            fn()
        eng.sync()
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        for _ in range(steps):
            fn()
        e1.record(stream)
        eng.sync()
        barrier()
        return reduce_max(e0.elapsed_time(e1))

    def time_host(fn, steps, warm):
        for _ in range(warm):
            fn()
        barrier()
        t0 = time.perf_counter()
        for _ in range(steps):
            fn()
        torch.cuda.synchronize()
        ms = reduce_max((time.perf_counter() - t0) * 1e3)
        barrier()
        return ms

    # ---- device-resident timing (headline `value`), per-launch CUDA events on for the roofline -------------------
    W = max(args.warmup, 3)
    for _ in range(W):
        step_dev()
    eng.sync()
    sampler = ClockSampler(local_rank)
    l0 = eng.launch_count
    eng.profile(True)
    barrier()
    sampler.start()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(stream)
    for _ in range(args.steps):
        step_dev()
    e1.record(stream)
    eng.sync()
    barrier()
    clocks = sampler.stop()
    ms_dev = reduce_max(e0.elapsed_time(e1))
    launches = eng.launch_count - l0
    prof = eng.profile_read()
    eng.profile(False)

    # ---- end to end through the C-ABI with host buffers ----------------------------------------------------------
    ms_e2e = time_host(step_host, args.steps, 2)
    u8_of_timed_step = out8_pin.numpy().reshape(args.nfeat, HW, HW, 3).copy() if rank == 0 and not args.no_extras else None

    # ---- D2H rate of this rank while every rank copies at once (the e2e limiter at N > 1) ------------------------
    def d2h_rate(src, dst):
        barrier()
        t0 = time.perf_counter()
        for _ in range(3):
            dst.copy_(src, non_blocking=True)
        torch.cuda.synchronize()
        dt = time.perf_counter() - t0
        t = torch.tensor([3 * src.numel() * src.element_size() / dt / 1e9], dtype=torch.float64, device="cuda")
        if dist is not None:
            dist.all_reduce(t, op=dist.ReduceOp.MIN)
        return float(t.item())
    d2h_u8, d2h_f32 = d2h_rate(out8_dev, out8_pin), d2h_rate(out_dev, out_pin)

    extras = {}
    if not args.no_extras:
        n_x = max(2, min(args.steps, 3))
        ms = time_dev(lambda: step_dev(F32), n_x, 1)
        ms_h = time_host(lambda: step_host(F32), n_x, 1)
        tot = args.nfeat * n_x * world
        extras["fp32_output"] = {
            "what": "the same step with the raw float32 projections as the result (what visualize_all_layers returns)",
            "value": tot / (ms / 1e3), "e2e": tot / (ms_h / 1e3), "unit": UNIT, "d2h_bytes_per_step": int(out_elems * 4),
            "e2e_limiter": "device-to-host copy of 308 MB per rank and step: %.1f GB/s per rank with all %d ranks copying"
                           % (d2h_f32, world)}
        if prec == "bf16":
            eng.profile(True)
            ms = time_dev(lambda: step_dev(U8, 0), n_x, 1)
            p32 = eng.profile_read()
            eng.profile(False)
            ms_h = time_host(lambda: step_host(U8, 0), n_x, 1)
            c = p32["conv_tc"]
            extras["fp32_mode"] = {
                "what": "the same step in the reference-precision mode (DCV_FP32: rel-L2 <= 1e-5): backward convs on tcgen05 with "
                        "operands scaled per projection and split into two fp16 planes, three products, fp32 register accumulation",
                "value": tot / (ms / 1e3), "e2e": tot / (ms_h / 1e3), "unit": UNIT, "ms_per_step": ms / n_x,
                "roofline": {"bound": "tensor", "kernel": "conv_tc (split-operand)", "achieved": c["flops"] / (c["ms"] / 1e3) / 1e12,
                             "peak": peaks()["tf_sus"], "unit": "TFLOP/s (algorithmic; the kernel issues 3 MMA flops per algorithmic flop)",
                             "frac": c["flops"] / (c["ms"] / 1e3) / 1e12 / peaks()["tf_sus"]} if c["ms"] > 0 else None}

    if rank == 0 and not args.no_extras:
        extras["parity"] = check_outputs(eng, li, img, prec, u8_of_timed_step, args)
    if not args.no_extras:
        extras["config4"] = bench_config4(eng.lib, model, dist, rank, world, local_rank, prec, barrier, reduce_max)
        extras["config5"] = bench_config5(dist, rank, world, local_rank, prec, barrier, reduce_max)

    if rank != 0:
        if dist is not None:
            dist.destroy_process_group()
        return
    pk = peaks()
    total = args.nfeat * args.steps * world
    value = total / (ms_dev / 1e3)
    e2e = total / (ms_e2e / 1e3)

    # ---- roofline of the dominant kernel (largest share of the step's device time) ----------------
    dom = max(prof, key=lambda k: prof[k]["ms"])
    d = prof[dom]
    share = d["ms"] / max(sum(v["ms"] for v in prof.values()), 1e-9)
    if dom in ("conv_fp32", "conv_tc", "rank1", "conv_fwd"):
        per_launch = d["flops"] / max(d["launches"], 1)
        achieved = d["flops"] / (d["ms"] / 1e3) / 1e12
        roof = {"bound": "tensor", "kernel": dom, "achieved": achieved, "peak": pk["tf_sus"], "unit": "TFLOP/s",
                "frac": achieved / pk["tf_sus"], "traffic": None, "peak_source": pk["source"] + " (sustained bf16)",
                "algorithmic_flops_per_launch": per_launch, "avg_launch_ms": d["ms"] / max(d["launches"], 1),
                "launches": d["launches"], "share_of_step": share}
    else:
        achieved = d["bytes"] / (d["ms"] / 1e3) / 1e9
        roof = {"bound": "hbm", "kernel": dom, "achieved": achieved, "peak": pk["hbm"], "unit": "GB/s",
                "frac": achieved / pk["hbm"], "traffic": None, "peak_source": pk["source"],
                "algorithmic_bytes_per_launch": d["bytes"] / max(d["launches"], 1),
                "avg_launch_ms": d["ms"] / max(d["launches"], 1), "launches": d["launches"], "share_of_step": share}
    # DRAM traffic per launch of the dominant class: NOT measured in this run (it needs ncu) — the figure of the
    # committed `ncu --set full`-class capture of this same step, named with the commit it was taken at
    for name in ("r2_traffic.json", "r1_traffic.json"):
        tpath = os.path.join(ROOT, "profiles", name)
        if os.path.exists(tpath) and args.nfeat == NFEAT and args.layer == LAYER and prec == "bf16":
            t = json.load(open(tpath)).get(dom)
            if t:
                roof["traffic"] = t["dram_bytes_per_launch"]
                roof["traffic_source"] = "committed ncu capture, not re-measured by this run: " + t["source"]
                roof["algorithmic_bytes_per_launch"] = d["bytes"] / max(d["launches"], 1)
                break
    classes = {k: {"ms_per_step": v["ms"] / args.steps, "launches_per_step": v["launches"] / args.steps,
                   "tflops": (v["flops"] / (v["ms"] / 1e3) / 1e12) if v["ms"] > 0 and v["flops"] else None,
                   "gbs": (v["bytes"] / (v["ms"] / 1e3) / 1e9) if v["ms"] > 0 and v["bytes"] else None}
               for k, v in prof.items() if v["launches"]}
    whole_flops = sum(v["flops"] for v in prof.values()) / args.steps
    roof["whole_step"] = {"algorithmic_tflop": whole_flops / 1e12, "achieved": whole_flops / (ms_dev / args.steps / 1e3) / 1e12,
                          "frac": whole_flops / (ms_dev / args.steps / 1e3) / 1e12 / pk["tf_sus"]}

    dream = None
    if not args.no_dream:
        dream = bench_dream(args, local_rank)

    cpu = None
    if not args.no_cpu_baseline:
        sample = cpu_sample_size(args.layer, args.nfeat, args.cpu_seconds)
        r, s, cores = cpu_projection_rate(args.layer, sample)
        cpu = {"value": r, "unit": UNIT, "cores": cores, "kind": "port",
               "sample": "forward to %s + %d of the %d projections + deprocess_image of each, vectorised torch-CPU restatement of "
                         "reference deepdream.py, %.1f s (TensorFlow 1.13 CPU path unavailable on this box)"
                         % (args.layer, sample, args.nfeat, s)}

    cfg = workload_config(args.layer, args.nfeat)
    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": W,
        "ms_per_step": ms_dev / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "bf16" if prec == "bf16" else "f32", "data": "synthetic",
        "config": cfg,
        "precision": "forward tape fp32-accurate on tcgen05 (3-way bf16 operand split, 6 products, ~1e-6 per layer), backward chain "
                     + ("bf16 on tcgen05 (rel-L2 <= 2e-2)" if prec == "bf16" else "fp32-accurate on tcgen05 (split operands, rel-L2 <= 1e-5)"),
        "timing": "CUDA events on the library's stream around the %d timed steps; per-launch event pairs are recorded inside the "
                  "timed region for the roofline (%d launches per step)" % (args.steps, launches // max(args.steps, 1)),
        "e2e": {"value": e2e, "unit": UNIT, "ms_per_step": ms_e2e / args.steps,
                "h2d_bytes_per_step": int(img.nbytes + feats.nbytes), "d2h_bytes_per_step": int(out_elems),
                "d2h_gbs_per_rank_all_ranks_copying": d2h_u8, "numa": numa},
        "gpu_launches": int(launches),
        "roofline": roof,
        "kernel_classes": classes,
        "cpu_baseline": cpu,
        "cpu_baseline_reference": reference_faithful_record(),
        "clocks": clocks,
        "deepdream": dream,
    }
    line.update(extras)
    print(json.dumps(line), flush=True)
    if dist is not None:
        dist.destroy_process_group()


def check_outputs(eng, li, img, prec, u8_host, args):
    """Outside the timed region: 8 of the timed step's outputs against the oracle.  The uint8 results must be
    deprocess_image of the library's float32 projections of the same maps (+-1 level), and those must match the
    oracle's projections within the mode's tolerance."""
    from oracle import keras_shim, restatement as rs
    pick = [int(f) for f in np.linspace(0, args.nfeat - 1, 8).astype(int)]
    eng.forward(img, li)
    # the whole step again with float32 results (same chunk schedule, hence the same kernels as the timed step)
    f32 = eng.project(li, np.arange(args.nfeat, dtype=np.int32), precision=prec)[pick, 0]
    tape = rs.FastTape(keras_shim.VGG16(include_top=False, input_shape=(HW, HW, 3)), args.layer, img)
    want = tape.project(pick)
    rel, lvl = 0.0, 0
    peak = max(float(np.abs(w).max()) for w in want)
    for k, f in enumerate(pick):
        if np.abs(want[k]).max() >= 0.05 * peak:          # almost-dead maps inherit the forward's last-ulp noise (tests/)
            rel = max(rel, float(np.linalg.norm(f32[k] - want[k]) / np.linalg.norm(want[k])))
        ref8 = rs.deprocess_image(f32[k].copy())
        lvl = max(lvl, int(np.abs(ref8.astype(np.int16) - u8_host[f].astype(np.int16)).max()))
    tol = 2e-2 if prec == "bf16" else 1e-5
    return {"maps_checked": pick, "max_rel_l2_vs_oracle": rel, "tolerance": tol, "max_uint8_level_difference": lvl,
            "ok": bool(rel < tol and lvl <= 1), "sha_of_step_output": __import__("hashlib").sha1(u8_host.tobytes()).hexdigest()}


def bench_config4(lib, model, dist, rank, world, local_rank, prec, barrier, reduce_max):
    """BASELINE configs[3]: 64 images x 512 block5_conv3 maps, sharded by image over the ranks through the product API
    (parallel.visualize_batch), uint8 results landing in host memory, no exchange."""
    import torch
    from deconv_api_b200 import parallel, synth, vgg16
    m = vgg16.VGG16(weights=None, include_top=False, input_shape=(HW, HW, 3))
    if dist is not None:
        parallel.broadcast_model_weights(m, dist, src=0, device=torch.device("cuda", local_rank))
    images = synth.image_batch(64, HW, HW, seed0=0)
    kw = dict(dist=dist, precision=prec, out_format="u8", images_per_tape=8, device=local_rank, sink=lambda i, f, a: None)
    parallel.visualize_batch(m, images, "block5_conv3", **kw)      # warm-up with the timed call's own block shapes: workspace, pinned sink buffer
    barrier()
    t0 = time.perf_counter()
    parallel.visualize_batch(m, images, "block5_conv3", **kw)
    torch.cuda.synchronize()
    s = reduce_max(time.perf_counter() - t0)
    barrier()
    tflop = 977.5
    return {"what": "BASELINE configs[3]: 64 images x 512 block5_conv3 feature maps, image-sharded over %d rank(s) with "
                    "parallel.visualize_batch, uint8 projections to host memory (4.9 GB), wall clock, max over ranks" % world,
            "seconds": s, "projections_per_s": 64 * 512 / s, "algorithmic_tflop": tflop,
            "tflops_per_gpu": tflop / s / world, "frac_of_sustained_bf16_peak": tflop / s / world / peaks()["tf_sus"]}


def bench_config5(dist, rank, world, local_rank, prec, barrier, reduce_max):
    """BASELINE configs[4]: DeepDream on 2048x2048, 4 octaves x 20 iterations, a batch of 8 images sharded over the ranks."""
    import torch
    from deconv_api_b200 import parallel, synth, vgg16
    m = vgg16.VGG16(weights=None, include_top=False, input_shape=(None, None, 3))
    if dist is not None:
        parallel.broadcast_model_weights(m, dist, src=0, device=torch.device("cuda", local_rank))
    S, n_img = 2048, 8
    base = synth.caffe_preprocess(synth.smooth_image_u8(0, S, S))
    images = np.concatenate([np.roll(base, 97 * i, axis=2) for i in range(n_img)], axis=0)
    kw = dict(dist=dist, device=local_rank, num_octave=4, iterations=20, precision=prec)
    parallel.dream_batch(m, images[:world], **kw)                                        # warm-up: buffers, graphs
    barrier()
    t0 = time.perf_counter()
    parallel.dream_batch(m, images, **kw)
    torch.cuda.synchronize()
    s = reduce_max(time.perf_counter() - t0)
    barrier()
    return {"what": "BASELINE configs[4]: DeepDream 2048x2048, 4 octaves x 20 iterations, 8 images sharded over %d rank(s) with "
                    "parallel.dream_batch (derived spec, no reference implementation), host image in / out, wall clock" % world,
            "seconds": s, "iterations_per_s": n_img * 80 / s, "algorithmic_tflop": 194.9 * n_img,
            "frac_of_sustained_bf16_peak": 194.9 * n_img / s / world / peaks()["tf_sus"]}


def bench_dream(args, local_rank):
    """Secondary BASELINE metric (configs[2]): DeepDream iterations/s on a SxS synthetic image, 3 octaves x 20
    iterations, mixed-layer loss (block3_conv3, block4_conv3, block5_conv3).  No reference implementation exists
    (derived spec); --precision bf16 = tcgen05 forward and backward on a bf16 tape."""
    import torch
    from deconv_api_b200 import dream as dm, synth, vgg16
    S = args.dream_size
    model = vgg16.VGG16(weights=None, include_top=False, input_shape=(None, None, 3))
    eng = model.bound_engine(local_rank, upto="block5_conv3")
    layers = dm.DEFAULT_LAYERS
    idx = [eng.index[n] for n in layers]
    coeffs = [layers[n] for n in layers]
    shapes = dm.octave_shapes(S, S, 3, 1.4)
    iters = 20
    jt = dm.jitter_offsets(3 * iters, 32, 0)
    img = synth.caffe_preprocess(synth.smooth_image_u8(0, S, S))
    img_dev = torch.from_numpy(img).cuda()
    out_dev = torch.empty_like(img_dev)
    stream = torch.cuda.ExternalStream(eng.stream, device=local_rank)
    eng.dream_dev(img_dev.data_ptr(), 1, S, S, idx, coeffs, shapes, iters, 1.0, jt, out_dev.data_ptr(), precision=args.precision)   # warm-up
    eng.sync()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    reps = 2
    l0 = eng.launch_count
    e0.record(stream)
    for _ in range(reps):
        eng.dream_dev(img_dev.data_ptr(), 1, S, S, idx, coeffs, shapes, iters, 1.0, jt, out_dev.data_ptr(), precision=args.precision)
    e1.record(stream)
    eng.sync()
    ms = e0.elapsed_time(e1) / reps
    t0 = time.perf_counter()
    eng.dream(img, idx, coeffs, shapes, iters, 1.0, jt, precision=args.precision)
    ms_e2e = (time.perf_counter() - t0) * 1e3
    fwd_gmac = sum(15.347 * (h * w) / (224.0 * 224.0) for h, w in shapes)          # conv stack to block5_conv3
    return {"metric": "deepdream_iters_per_sec", "value": 3 * iters / (ms / 1e3), "unit": "iterations/s",
            "e2e_value": 3 * iters / (ms_e2e / 1e3), "ms_per_dream": ms, "dtype": "bf16" if args.precision == "bf16" else "f32",
            "config": {"workload": "DeepDream %dx%d, octaves %s x %d iterations, loss on block3_conv3/block4_conv3/block5_conv3 "
                                   "(BASELINE configs[2]; derived spec, no reference implementation)" % (S, S, shapes, iters)},
            "tflops": 2 * 2 * fwd_gmac * iters / (ms / 1e3) / 1e3, "gpu_launches": (eng.launch_count - l0) // reps}


def main():
    args = parse()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if args.impl == "reference":
        run_reference(args, rank, world)
    else:
        run_ours(args, rank, world, local_rank)


if __name__ == "__main__":
    main()
