#!/usr/bin/env python
"""bench.py — VGG16 deconvnet projections/s on B200 (BASELINE.json metric), one JSON line.

A "step" is one pass of the hot path over one batch of synthetic input.  Workload at every N
(weak scaling, no data-path collective): each rank takes its own 224x224 synthetic image, runs the
forward tape to block4_conv3 once and projects ALL 512 feature maps of block4_conv3 to pixel space
(BASELINE.json configs[1]).  `value` is timed with inputs and outputs resident in HBM; `e2e` is the
same step through the C-ABI with HOST buffers (pinned image in, 512 projections out).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--precision bf16|fp32]

N > 1: launched by torchrun, one rank per GPU; weights are broadcast once over NCCL, then no
collective touches the data path.  --impl reference times the CPU port of the reference (the oracle;
TensorFlow 1.13 is not installable here) on the box's host cores, rank 0 only.
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


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--precision", default=os.environ.get("DCV_BENCH_PRECISION", "bf16"), choices=["bf16", "fp32"])
    ap.add_argument("--layer", default=LAYER)
    ap.add_argument("--nfeat", type=int, default=NFEAT)
    ap.add_argument("--cpu-sample", type=int, default=512, help="projections in the cpu_baseline sample (~10 s of CPU work)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-dream", action="store_true", help="skip the secondary DeepDream measurement")
    ap.add_argument("--dream-size", type=int, default=512)
    return ap.parse_args()


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
def cpu_projection_rate(layer, nproj, seed=0, batch=16, repeats=1):
    """Times the oracle port (vectorised torch-CPU restatement of reference deepdream.py) on the host:
    forward tape + `nproj` projections of `layer`, all host threads.  Returns (proj/s, seconds, cores)."""
    import torch
    from oracle import keras_shim, restatement as rs
    from deconv_api_b200 import synth
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    model = keras_shim.VGG16(include_top=False, input_shape=(HW, HW, 3))
    img = synth.image_batch(1, HW, HW, seed)
    best = None
    for _ in range(repeats):
        t0 = time.perf_counter()
        tape = rs.FastTape(model, layer, img)
        feats = list(range(nproj))
        for done in range(0, nproj, batch):
            tape.project(feats[done:done + batch])
        dt = time.perf_counter() - t0
        best = dt if best is None else min(best, dt)
    return nproj / best, best, cores


def run_reference(args, rank, world):
    if rank != 0:
        return
    sample = max(8, min(args.cpu_sample, args.nfeat, 128))     # ~2-3 s of CPU work per step
    for _ in range(min(args.warmup, 1)):
        cpu_projection_rate(args.layer, 8)
    rates, secs = [], []
    for _ in range(args.steps):
        r, s, cores = cpu_projection_rate(args.layer, sample)
        rates.append(r); secs.append(s)
    value = sample * len(secs) / sum(secs)
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": 1e3 * sum(secs) / len(secs), "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": "VGG16 224x224, forward to %s + projections of its feature maps, batch 1 "
                               "(BASELINE configs[1]); each step = forward + %d projections (bounded sample of the 512)"
                               % (args.layer, sample)},
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": cores, "kind": "port",
                         "sample": "forward to %s + %d projections per step, vectorised torch-CPU restatement of reference "
                                   "deepdream.py (TensorFlow 1.13 CPU path unavailable on this box)" % (args.layer, sample)},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


# --------------------------------------------------------------------------------------------- GPU arm
def run_ours(args, rank, world, local_rank):
    import torch
    from deconv_api_b200 import synth, vgg16
    torch.cuda.set_device(local_rank)
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

    img = synth.image_batch(1, HW, HW, seed0=rank)                    # this rank's unit of work
    out_elems = args.nfeat * HW * HW * 3
    img_dev = torch.from_numpy(img).cuda()
    out_dev = torch.empty(out_elems, dtype=torch.float32, device="cuda")
    img_pin = torch.from_numpy(img).pin_memory()
    out_pin = torch.empty(out_elems, dtype=torch.float32).pin_memory()
    stream = torch.cuda.ExternalStream(eng.stream, device=local_rank)

    def step_dev():
        eng.forward_dev(img_dev.data_ptr(), 1, HW, HW, li)
        eng.project_dev(li, feats, out_dev.data_ptr(), precision=prec)

    def step_host():
        eng._ck(eng.lib.dcv_forward(eng.ctx, ctypes.c_void_p(img_pin.data_ptr()), 1, HW, HW, li))
        eng._ck(eng.lib.dcv_project(eng.ctx, li, feats.ctypes.data_as(ctypes.c_void_p), args.nfeat, 0,
                                    1 if prec == "bf16" else 0, ctypes.c_void_p(out_pin.data_ptr())))

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

    # ---- device-resident timing ------------------------------------------------------------------
    for _ in range(max(args.warmup, 3)):
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

    # ---- end to end through the C-ABI with host buffers ------------------------------------------
    for _ in range(2):
        step_host()
    barrier()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        step_host()
    torch.cuda.synchronize()
    ms_e2e = reduce_max((time.perf_counter() - t0) * 1e3)
    barrier()

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
    # DRAM traffic per launch of the dominant class from the committed ncu capture of this same step
    tpath = os.path.join(ROOT, "profiles", "r1_traffic.json")
    if os.path.exists(tpath) and args.nfeat == NFEAT and args.layer == LAYER and prec == "bf16":
        t = json.load(open(tpath)).get(dom)
        if t:
            roof["traffic"] = t["dram_bytes_per_launch"]
            roof["traffic_source"] = t["source"]
            roof["algorithmic_bytes_per_launch"] = d["bytes"] / max(d["launches"], 1)
    classes = {k: {"ms_per_step": v["ms"] / args.steps, "launches_per_step": v["launches"] / args.steps,
                   "tflops": (v["flops"] / (v["ms"] / 1e3) / 1e12) if v["ms"] > 0 and v["flops"] else None,
                   "gbs": (v["bytes"] / (v["ms"] / 1e3) / 1e9) if v["ms"] > 0 and v["bytes"] else None}
               for k, v in prof.items() if v["launches"]}

    dream = None
    if not args.no_dream:
        dream = bench_dream(args, local_rank)

    cpu = None
    if not args.no_cpu_baseline:
        r, s, cores = cpu_projection_rate(args.layer, min(args.cpu_sample, args.nfeat))
        cpu = {"value": r, "unit": UNIT, "cores": cores, "kind": "port",
               "sample": "forward to %s + %d of the %d projections, vectorised torch-CPU restatement of reference "
                         "deepdream.py, %.1f s (TensorFlow 1.13 CPU path unavailable on this box)"
                         % (args.layer, min(args.cpu_sample, args.nfeat), args.nfeat, s)}

    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": max(args.warmup, 3),
        "ms_per_step": ms_dev / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "bf16" if prec == "bf16" else "f32", "data": "synthetic",
        "config": {"workload": "VGG16 224x224 random-init, forward to %s (fp32) + projection of all %d feature maps to pixel "
                               "space, batch 1 per GPU (BASELINE configs[1]); one image per rank" % (args.layer, args.nfeat),
                   "precision": "forward tape fp32-accurate on tcgen05 (3-way bf16 operand split, 6 products, ~1e-6 per layer), "
                                "backward chain %s" % ("bf16 on tcgen05" if prec == "bf16" else "fp32 FFMA"),
                   "l2": "no flush needed: each step streams >6 GB of intermediates, far above the 126 MB L2",
                   "parallelism": "replicated weights, image-sharded, no per-step collective"},
        "e2e": {"value": e2e, "unit": UNIT, "ms_per_step": ms_e2e / args.steps,
                "h2d_bytes_per_step": int(img.nbytes + feats.nbytes), "d2h_bytes_per_step": int(out_elems * 4)},
        "gpu_launches": int(launches),
        "roofline": roof,
        "kernel_classes": classes,
        "cpu_baseline": cpu,
        "clocks": clocks,
        "deepdream": dream,
    }
    print(json.dumps(line), flush=True)
    if dist is not None:
        dist.destroy_process_group()


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
