#!/usr/bin/env python
"""Headline benchmark: NewCRFs Swin-L (``large07``) inference images/s at
352x1216, bf16 autocast, batch 8 per GPU, synthetic KITTI-shape input
(BASELINE.json configs[1]).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--mode infer|train]

One JSON line on rank 0.  ``value`` = device-resident throughput, ``e2e`` = the
same through the serving loop ``harness.infer_host_batches`` with pinned-host
images copied in and the depth map copied out every step.  ``roofline`` is for
the window-attention forward kernel (all its launches inside the timed region,
CUDA events on the launching stream).  Extra keys of the default line:

* ``train``: BASELINE configs[2] -- fwd + SILog + bwd + clip + AdamW at 480x640, batch 8/GPU, DDP over NCCL at
  N > 1 (the step, with its gradient all-reduce, replayed from a CUDA graph; the eager step beside it);
* ``reference_gpu_eager`` (N = 1): the UNMODIFIED reference ``NewCRFDepth`` (``oracle/_ref``) on the same B200,
  fp32 and ``torch.autocast(bf16)``, same batch, CUDA events;
* ``cpu_baseline`` (N = 1): the same unmodified reference modules on the host cores.

``--impl reference`` is the reference arm: the unmodified reference on the host cores only; it imports nothing
from the product package (no product ``.so`` is loaded in that process).
"""
from __future__ import annotations

import argparse
import ctypes
import gc
import json
import os
import statistics
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402

VERSION, HEIGHT, WIDTH, BATCH, MAX_DEPTH = "large07", 352, 1216, 8, 80.0
METRIC = "NewCRFs Swin-L images/sec @352x1216 (bf16 inference, batch 8/GPU)"
TRAIN_METRIC = "NewCRFs Swin-L training images/sec @480x640 (bf16 autocast fwd + SILog + bwd + clip + AdamW, batch 8/GPU)"
TRAIN_LEG_TIMEOUT_S = 420.0


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--mode", default="infer", choices=["infer", "train"],
                    help="infer: BASELINE configs[1] (the headline metric) with configs[2] as the `train` extra key; "
                         "train: configs[2] as the line itself (fwd+SILog+bwd+AdamW at 480x640)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-graph", action="store_true", help="launch eagerly everywhere; no CUDA graph replay")
    ap.add_argument("--graph-ddp", action="store_true", help="(the default now; kept for compatibility)")
    ap.add_argument("--no-graph-ddp", action="store_true",
                    help="training at N > 1: keep the eager DDP step; do not capture it (with its NCCL all-reduce) in a CUDA graph")
    ap.add_argument("--no-train-leg", action="store_true", help="default mode: skip the `train` extra key (BASELINE configs[2])")
    ap.add_argument("--no-reference-gpu", action="store_true", help="skip the `reference_gpu_eager` extra key")
    ap.add_argument("--no-graph-timing", action="store_true",
                    help="skip the instrumented CUDA-graph replays behind `roofline.graph_replay`")
    ap.add_argument("--bf16-grad-hook", action="store_true",
                    help="training at N > 1: DDP's bf16_compress_hook on the gradient buckets (halves the all-reduce bytes)")
    ap.add_argument("--version", default=VERSION)
    ap.add_argument("--batch", type=int, default=BATCH)
    return ap.parse_args()


def dist_env():
    return dict(rank=int(os.environ.get("RANK", 0)), local_rank=int(os.environ.get("LOCAL_RANK", 0)),
                world=int(os.environ.get("WORLD_SIZE", 1)))


def peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        d = json.load(open(path))
        return d["hbm_gbs"], d["bf16_tflops_sustained"], "measured"
    return 6650.0, 1400.0, "fallback"  # B200_PROFILING.md fallback


class ClockSampler(threading.Thread):
    """SM clock + throttle reasons during the timed region (pynvml)."""

    def __init__(self, index):
        super().__init__(daemon=True)
        self.index, self.samples, self.reasons, self.max_mhz, self.stop_flag = index, [], set(), None, False

    def run(self):
        try:
            import pynvml as nv
            nv.nvmlInit()
            h = nv.nvmlDeviceGetHandleByIndex(self.index)
            self.max_mhz = nv.nvmlDeviceGetMaxClockInfo(h, nv.NVML_CLOCK_SM)
            names = {
                nv.nvmlClocksThrottleReasonHwSlowdown: "hw_slowdown",
                nv.nvmlClocksThrottleReasonHwThermalSlowdown: "hw_thermal_slowdown",
                nv.nvmlClocksThrottleReasonSwThermalSlowdown: "sw_thermal_slowdown",
                nv.nvmlClocksThrottleReasonSwPowerCap: "sw_power_cap",
            }
            while not self.stop_flag:
                self.samples.append(nv.nvmlDeviceGetClockInfo(h, nv.NVML_CLOCK_SM))
                r = nv.nvmlDeviceGetCurrentClocksThrottleReasons(h)
                self.reasons |= {n for bit, n in names.items() if r & bit}
                time.sleep(0.1)
        except Exception as e:  # clocks are a report, not a dependency
            self.reasons.add(f"unavailable:{type(e).__name__}")

    def stop(self):
        self.stop_flag = True
        if self.is_alive():
            self.join(timeout=2)

    def summary(self):
        return {"sm_mhz": statistics.median(self.samples) if self.samples else None,
                "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons)}


# ---- the reference's own code on the host cores ------------------------------------------------------------------

def cpu_reference(version, steps, warmup, threads):
    """Times the UNMODIFIED reference ``NewCRFDepth(version)`` (oracle/_ref or /root/reference through the shim), fp32,
    ``torch.no_grad()``, one 352x1216 image per step, on ``threads`` host threads.  Falls back to the oracle port (the
    product's module graph with the oracle attention, ``kind: "port"``) only when no reference tree is staged."""
    from oracle import ref_model
    torch.set_num_threads(threads)
    if ref_model.available():
        from oracle import reference_shim
        model = ref_model.build_newcrfs(version, MAX_DEPTH, seed=0)
        kind, src = "reference", f"unmodified model/NewCRFs/NewCRFDepth.py from {reference_shim.source()} (timm / mmcv: 4-symbol shim)"
    else:
        from oracle.cpu_model import build_newcrfs   # imports the product package: last resort, labelled as a port
        model = build_newcrfs(version, seed=0, max_depth=MAX_DEPTH)
        kind, src = "port", "oracle/cpu_model.py (oracle/_ref not staged: run __graft_entry__.build() where /root/reference exists)"
    img = ref_model.images(1, HEIGHT, WIDTH, seed=0)
    times = ref_model.time_forward(lambda: model(img), steps, warmup)
    return {"value": steps / sum(times), "unit": "images/s", "cores": threads, "kind": kind, "source": src,
            "sample": f"1 image (of the batch-8 workload) per step, fp32, torch CPU, {warmup} warm-up + {steps} timed forwards",
            "median_images_per_s": 1.0 / statistics.median(times), "best_images_per_s": 1.0 / min(times),
            "worst_images_per_s": 1.0 / max(times), "ms_per_step": sum(times) / steps * 1e3}


def cpu_reference_configs0(threads):
    """BASELINE configs[0]: the NeWCRF decoder block (crf3 .. crf0, window 7) of the reference ``tiny07`` model on its own
    encoder + PSP features of one synthetic 352x704 image, fp32 on the CPU (model/NewCRFs/NewCRFDepth.py:131-137):
    2 warm-ups, median / best of 5 on all threads, median of 3 on one thread."""
    from oracle import ref_model
    if not ref_model.available():
        return None
    model = ref_model.build_newcrfs("tiny07", MAX_DEPTH, seed=0)
    img = ref_model.images(1, 352, 704, seed=0)
    with torch.no_grad():
        feats = model.backbone(img)
        ppm = model.decoder(feats)
    out = {"workload": "NeWCRF decoder crf3..crf0 (window 7) on tiny07 encoder + PSP features, 1x3x352x704, fp32, CPU"}
    for label, n_threads, steps, warmup in (("all_threads", threads, 5, 2), ("one_thread", 1, 3, 1)):
        torch.set_num_threads(n_threads)
        t = ref_model.time_forward(lambda: ref_model.decoder_only_forward(model, feats, ppm), steps, warmup)
        out[label] = {"threads": n_threads, "median_ms": statistics.median(t) * 1e3, "best_ms": min(t) * 1e3, "n": steps}
    torch.set_num_threads(threads)
    return out


def run_reference(args, rank):
    """Reference arm: the reference's own CPU implementation, every host thread, one image of the workload per step.
    Imports nothing from the product package."""
    if rank != 0:
        return
    cores = os.cpu_count() or 1
    base = cpu_reference(args.version, args.steps, args.warmup, cores)
    line = {
        "impl": "reference", "metric": METRIC, "value": base["value"], "unit": "images/s", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": base["ms_per_step"], "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": f"NewCRFs {args.version} full-model inference, 352x1216 KITTI shape",
                   "sample": "1 image per step on the host CPU"},
        "cpu_baseline": base,
        "e2e": {"value": base["value"], "unit": "images/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    try:
        line["configs0"] = cpu_reference_configs0(cores)
    except Exception as e:   # an extra key must never cost the line
        line["configs0"] = {"error": f"{type(e).__name__}: {e}"}
    emit(line)


def reference_gpu_eager(args, dev, flush):
    """The unmodified reference modules on the B200 (PyTorch eager: ATen / cuBLAS / cuDNN), same batch as `value`,
    fp32 (TF32 off, torch's default) and torch.autocast(bf16) -- the "existing Blackwell implementation" (SURVEY F8)."""
    from oracle import ref_model, reference_shim
    if not ref_model.available():
        return {"unavailable": "oracle/_ref not staged"}
    model = ref_model.build_newcrfs(args.version, MAX_DEPTH, seed=0, device=dev)
    x = ref_model.images(args.batch, HEIGHT, WIDTH, seed=0).to(dev)
    out = {"source": reference_shim.source(), "batch": args.batch, "timing": "CUDA events per step, L2 flushed, 2 warm-ups + 5 steps"}
    for label, bf16 in (("bf16_autocast", True), ("fp32", False)):
        def step():
            with torch.autocast("cuda", dtype=torch.bfloat16, enabled=bf16):
                model(x)
        ms = ref_model.time_forward_cuda(step, 5, 2, flush)
        out[label] = {"images_per_s": args.batch / (statistics.median(ms) * 1e-3), "ms_per_step": statistics.median(ms),
                      "best_ms": min(ms)}
    del model, x
    gc.collect()
    torch.cuda.empty_cache()
    return out


# ---- helpers of the product arm -------------------------------------------------------------------------------------

def roofline_of(recs, name, steps, t_res, traffic):
    hbm, tf, which = peaks()
    k_ms = sum(r[0] for r in recs)
    if not recs or not k_ms:
        return None
    k_bytes, k_flops = sum(r[1] for r in recs), sum(r[2] for r in recs)
    achieved = k_bytes / (k_ms * 1e-3) / 1e9
    by_shape = {}
    for r in recs:   # tag = ("fwd"|"bwd", window, shift, heads, H, W)
        by_shape.setdefault(f"ws{r[3][1]}_nH{r[3][3]}_{r[3][4]}x{r[3][5]}", []).append(r)
    return {"bound": "hbm", "kernel": name, "achieved": achieved, "peak": hbm, "unit": "GB/s", "frac": achieved / hbm,
            "peak_source": f"{which} (MEASURED_PEAKS.json hbm_gbs)",
            "traffic": (traffic.get(name.split()[1]) or {}).get("dram_bytes_per_launch"),
            "traffic_of": traffic.get(name.split()[1]),
            "launches": len(recs), "_steps": steps, "kernel_ms_per_step": k_ms / steps, "share_of_step": k_ms / 1e3 / t_res,
            "useful_tflops": k_flops / (k_ms * 1e-3) / 1e12,
            "tensor_frac_of_sustained_bf16": k_flops / (k_ms * 1e-3) / 1e12 / tf,
            "by_shape": {k: {"launches": len(v), "us_per_launch": sum(r[0] for r in v) / len(v) * 1e3,
                             "frac": sum(r[1] for r in v) / (sum(r[0] for r in v) * 1e-3) / 1e9 / hbm}
                         for k, v in sorted(by_shape.items())}}


def timed_graph_replays(ctx, build_graph, replay, steps):
    """Launch times INSIDE a CUDA-graph replay: `build_graph()` captures an instrumented copy of the graph (every
    attention / linear launch bracketed by two external event-record nodes, ops.capture_graph_timing), which is replayed
    `steps` times with an L2 flush before each; returns (records of all replays, seconds of the replays)."""
    from mde_b200 import ops
    ops.capture_graph_timing(True)
    try:
        g = build_graph()
    finally:
        ops.capture_graph_timing(False)
    recs, total = [], 0.0
    try:
        for i in range(steps + 1):                       # the first replay is a warm-up
            ctx.flush.zero_()
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record()
            replay(g)
            b.record()
            torch.cuda.synchronize()
            if i:
                recs += [r for r in ops.read_graph_timing() if r[0] >= 0]
                total += a.elapsed_time(b) / 1e3
    finally:
        del g
        gc.collect()
        ops.clear_graph_timing()
    return recs, total


def with_bracket_probe(roof, with_memset):
    """Adds what the event bracket costs around an EMPTY kernel (measured now, same stream; the backward's bracket also
    holds its memset) and the fraction net of that fixed cost.  `frac` stays the plain event-timed figure."""
    if roof is None:
        return roof
    from mde_b200 import _lib
    us = _lib.load().swin_attn_probe_empty_bracket_us(ctypes.c_void_p(torch.cuda.current_stream().cuda_stream), int(with_memset), 20)
    if us > 0:
        net_ms = roof["kernel_ms_per_step"] - us * 1e-3 * roof["launches"] / roof["_steps"]
        roof["empty_kernel_bracket_us"] = us
        roof["empty_kernel_bracket_note"] = ("median event bracket around an EMPTY 148x384 kernel" + (" preceded by the backward's memset" if with_memset else "")
                                             + " that follows a running kernel (swin_attn_probe_empty_bracket_us, measured in this run): "
                                             "event records, launch latency and completion -- the fixed cost inside every us_per_launch")
        if net_ms > 0:
            roof["frac_net_of_empty_bracket"] = roof["frac"] * roof["kernel_ms_per_step"] / net_ms
    roof.pop("_steps", None)
    return roof


def with_bracket_probe_of(recs, with_memset, name, steps, t_res, traffic):
    return with_bracket_probe(roofline_of(recs, name, steps, t_res, traffic), with_memset)


def with_graph_timing(roof, recs, name, steps, t_graph, traffic):
    """Adds the same roofline, timed inside the CUDA-graph replay, under `graph_replay` of an event-timed roofline."""
    if roof is None or not recs:
        return roof
    g = roofline_of(recs, name, steps, t_graph, traffic)
    if g:
        g.pop("_steps", None)
        roof["graph_replay"] = {k: g[k] for k in ("achieved", "frac", "launches", "kernel_ms_per_step", "share_of_step", "by_shape")}
        roof["graph_replay"]["timed"] = ("external event-record nodes around each launch inside an instrumented copy of the step's "
                                         "CUDA graph (no host launch gaps between the events and the kernel)")
    return roof


def roofline_linear_of(recs, steps, t_res):
    hbm, tf, which = peaks()
    if not recs:
        return None
    l_ms, l_bytes, l_flops = (sum(r[i] for r in recs) for i in range(3))
    out = {"bound": "tensor", "kernel": "tc_linear_kernel (per-token Linears of the blocks as tcgen05 GEMMs with fused epilogues; "
                                        "CTA pairs / cta_group::2 on the large-K shapes)",
           "achieved": l_flops / (l_ms * 1e-3) / 1e12, "peak": tf, "unit": "TFLOP/s", "frac": l_flops / (l_ms * 1e-3) / 1e12 / tf,
           "peak_source": f"{which} (MEASURED_PEAKS.json dense bf16)", "traffic": None,
           "algorithmic_GBps": l_bytes / (l_ms * 1e-3) / 1e9, "hbm_frac": l_bytes / (l_ms * 1e-3) / 1e9 / hbm,
           "launches": len(recs), "kernel_ms_per_step": l_ms / steps, "share_of_step": l_ms / 1e3 / t_res, "by_shape": {}}
    shapes = {}
    for r in recs:                                   # tag = ("lin", in_features, activation, out_features, rows, 0)
        shapes.setdefault((r[3][4], r[3][1], r[3][3], r[3][2]), []).append(r)
    for (rows, k, n, act), rs in sorted(shapes.items(), reverse=True):
        ms = sum(r[0] for r in rs)
        out["by_shape"][f"{rows}x{k}x{n}" + ("" if act == 1 else f"_epi{act}")] = {
            "launches": len(rs), "us_per_launch": ms / len(rs) * 1e3, "TFLOPs": sum(r[2] for r in rs) / (ms * 1e-3) / 1e12,
            "GBps": sum(r[1] for r in rs) / (ms * 1e-3) / 1e9}
    return out


def load_traffic():
    tpath = os.path.join(ROOT, "profiles", "roofline_traffic.json")
    return json.load(open(tpath)) if os.path.exists(tpath) else {}


class Ctx:
    """Per-process state shared by the legs."""

    def __init__(self, args, env):
        self.args, self.env = args, env
        self.rank, self.world, self.local = env["rank"], env["world"], env["local_rank"]
        self.dev = torch.device("cuda", self.local)
        torch.cuda.set_device(self.dev)
        self.flush = torch.empty(256 << 20, dtype=torch.uint8, device=self.dev)  # > 126 MB L2

    def barrier(self):
        if self.world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed_steps(self, step_fn, steps):
        """K steps, each bracketed by its own CUDA events with an L2 flush before it."""
        evs = []
        for _ in range(steps):
            self.flush.zero_()
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record()
            step_fn()
            b.record()
            evs.append((a, b))
        torch.cuda.synchronize()
        return sum(a.elapsed_time(b) for a, b in evs) / 1e3

    def max_over_ranks(self, values):
        t = torch.tensor([v or 0.0 for v in values], dtype=torch.float64, device=self.dev)
        if self.world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return t.tolist()


def infer_leg(ctx):
    """BASELINE configs[1].  Returns the pieces of the line (rank 0 uses them)."""
    from mde_b200 import _lib, ops
    from mde_b200.harness import GraphedInference, build_model, infer, infer_host_batches, prepare_inference, synthetic_images
    args, dev = ctx.args, ctx.dev
    model = build_model(args.version, MAX_DEPTH, seed=0, device=dev)
    host = synthetic_images(args.batch, HEIGHT, WIDTH, seed=ctx.rank, pin=True)
    images = host.to(dev)
    prepare_inference(model)   # eval, no grads, GEMM / conv weights stored in bf16 (what autocast reads anyway)
    out_host = torch.empty(args.batch, 1, HEIGHT, WIDTH).pin_memory()
    out_hosts = [out_host, torch.empty_like(out_host).pin_memory()]
    graphed = None if args.no_graph else GraphedInference(model, images)

    def step_resident():
        infer(model, images)

    def region_e2e(steps, g):
        """K steps through the serving loop (harness.infer_host_batches): every step's images come from pinned host
        memory and its depth map goes back to pinned host memory inside the timed region; the copies of neighbouring
        steps overlap the compute on copy streams.  One event pair around the region, L2 flushed before each forward."""
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize()
        a.record()
        infer_host_batches(model, (host for _ in range(steps)), [out_hosts[i & 1] for i in range(steps)],
                           before_step=ctx.flush.zero_, graphed=g)
        b.record()
        torch.cuda.synchronize()
        return a.elapsed_time(b) / 1e3

    for _ in range(args.warmup):
        step_resident()
        region_e2e(1, graphed)
    sampler = ClockSampler(ctx.local)
    records = []
    ctx.barrier()
    if os.environ.get("BENCH_NO_CLOCK_SAMPLER") != "1":   # diagnosis only: the JSON line then carries no clock samples
        sampler.start()
    launches0 = _lib.launch_count()
    ops.set_kernel_timing(records)
    t_res = ctx.timed_steps(step_resident, args.steps)
    ops.set_kernel_timing(None)
    launches = _lib.launch_count() - launches0
    ctx.barrier()
    # One event pair brackets the whole K-step serving region, so a single host hiccup (GC, the clock sampler taking
    # the GIL, a page fault in a pinned buffer) while the ~100 ms region is being enqueued lands in the number in full.
    # The region is therefore timed five times, each with exactly K steps and a synchronize on both sides, and the
    # median is reported.
    e2e_regions = sorted(region_e2e(args.steps, graphed) for _ in range(5))
    t_e2e = e2e_regions[2]
    t_e2e_eager = None
    if graphed is not None:   # the same serving loop with the forward launched eagerly: comparable with `value`
        t_e2e_eager = sorted(region_e2e(args.steps, None) for _ in range(3))[1]
    t_graph = None
    if graphed is not None:
        ctx.barrier()
        t_graph = ctx.timed_steps(lambda: graphed(images), args.steps)   # device-resident, the forward replayed as one graph
    ctx.barrier()
    sampler.stop()
    graph_records, t_graph_timed = [], None
    if graphed is not None and ctx.rank == 0 and not args.no_graph_timing:
        try:
            graph_records, t_graph_timed = timed_graph_replays(ctx, lambda: GraphedInference(model, images), lambda g: g(images), args.steps)
        except Exception as e:
            print(f"[bench] graph-replay launch timing skipped: {type(e).__name__}: {e}"[:300], file=sys.stderr)
    ctx.barrier()
    t_res, t_e2e, t_graph, t_e2e_eager = ctx.max_over_ranks([t_res, t_e2e, t_graph, t_e2e_eager])
    n_img = args.batch * ctx.world * args.steps
    res = dict(t_res=t_res, t_e2e=t_e2e, t_graph=t_graph or None, t_e2e_eager=t_e2e_eager or None, records=records,
               launches=int(launches), clocks=sampler.summary(), h2d=host.numel() * 4, d2h=out_host.numel() * 4,
               e2e_regions=[n_img / t for t in e2e_regions], n_img=n_img, graph_records=graph_records, t_graph_timed=t_graph_timed)
    del graphed, model, images
    gc.collect()
    torch.cuda.empty_cache()
    return res


def train_leg(ctx, steps, warmup):
    """BASELINE configs[2]: fwd + SILog + bwd + clip + AdamW at 480x640, batch 8/GPU, stock DDP over NCCL at N > 1.
    The eager step is timed first (its attention launches carry the roofline events), then the whole step -- with the
    bucketed gradient all-reduce at N > 1 -- is captured as one CUDA graph (harness.GraphedTrainStep) and timed again."""
    from mde_b200 import _lib, ops
    from mde_b200.harness import GraphedTrainStep, Trainer, build_model, replicas_in_sync, synthetic_depth, synthetic_images
    args, dev, world = ctx.args, ctx.dev, ctx.world
    height, width, max_depth = 480, 640, 10.0
    graph_step = not args.no_graph and (world == 1 or not args.no_graph_ddp)
    if graph_step and world > 1:
        # the DDP step is captured in a CUDA graph: torch's recipe wants the wrapper built, and the job stepped, on one
        # non-default stream
        torch.cuda.set_stream(torch.cuda.Stream(dev))
    model = build_model(args.version, max_depth, seed=0, device=dev)
    model.train()
    host = synthetic_images(args.batch, height, width, seed=ctx.rank, pin=True)
    images = host.to(dev)
    trainer = Trainer(model, ddp=world > 1, device_ids=[ctx.local] if world > 1 else None, capturable=graph_step)
    if args.bf16_grad_hook and world > 1:
        from torch.distributed.algorithms.ddp_comm_hooks import default_hooks
        trainer.model.register_comm_hook(None, default_hooks.bf16_compress_hook)
    gt_host = synthetic_depth(args.batch, height, width, max_depth, seed=100 + ctx.rank).pin_memory()
    gt = gt_host.to(dev)
    out_host = torch.empty(1).pin_memory()

    def step_resident():
        trainer.step(images, gt)

    def step_e2e_eager():
        x, y = host.to(dev, non_blocking=True), gt_host.to(dev, non_blocking=True)
        out_host.copy_(trainer.step(x, y).reshape(1), non_blocking=True)

    for _ in range(warmup):
        step_resident()
        step_e2e_eager()
    sampler = ClockSampler(ctx.local)
    records = []
    ctx.barrier()
    if os.environ.get("BENCH_NO_CLOCK_SAMPLER") != "1":
        sampler.start()
    launches0 = _lib.launch_count()
    ops.set_kernel_timing(records)
    t_res = ctx.timed_steps(step_resident, steps)
    ops.set_kernel_timing(None)
    launches = _lib.launch_count() - launches0
    ctx.barrier()
    gstep, graph_error, in_sync, t_graph = None, None, None, None
    step_e2e = step_e2e_eager
    if graph_step:
        try:
            gstep = GraphedTrainStep(trainer, images, gt, warmup=2)

            def step_e2e():   # noqa: F811  (the end-to-end step through the graph: H2D -> static inputs -> replay -> D2H)
                x, y = host.to(dev, non_blocking=True), gt_host.to(dev, non_blocking=True)
                out_host.copy_(gstep(x, y).reshape(1), non_blocking=True)
            for _ in range(2):
                step_e2e()
        except Exception as e:   # keep the eager numbers; say why the replay is missing
            gstep, graph_error, step_e2e = None, f"{type(e).__name__}: {e}"[:300], step_e2e_eager
    t_e2e = ctx.timed_steps(step_e2e, steps)
    if gstep is not None:
        t_graph = ctx.timed_steps(lambda: gstep(images, gt), steps)  # device-resident, the training step as one graph
        if world > 1:
            # every rank trains on its own images: the replicas stay identical only if the captured step really
            # all-reduces the gradients
            in_sync = replicas_in_sync(model)
    ctx.barrier()
    sampler.stop()
    graph_records, t_graph_timed = [], None
    if gstep is not None and world == 1 and not args.no_graph_timing:   # (one capture of the DDP step per process is enough risk)
        try:
            graph_records, t_graph_timed = timed_graph_replays(ctx, lambda: GraphedTrainStep(trainer, images, gt, warmup=2),
                                                               lambda g: g(images, gt), steps)
        except Exception as e:
            print(f"[bench] graph-replay launch timing (train) skipped: {type(e).__name__}: {e}"[:300], file=sys.stderr)
    t_res, t_e2e, t_graph = ctx.max_over_ranks([t_res, t_e2e, t_graph])
    res = dict(t_res=t_res, t_e2e=t_e2e, t_graph=t_graph or None, records=records, launches=int(launches),
               clocks=sampler.summary(), h2d=host.numel() * 4 + gt_host.numel() * 4, d2h=4, in_sync=in_sync,
               graph_error=graph_error, graphed=gstep is not None, n_img=args.batch * world * steps, steps=steps,
               bf16_grad_hook=bool(args.bf16_grad_hook and world > 1), graph_records=graph_records, t_graph_timed=t_graph_timed)
    del gstep, trainer, model
    gc.collect()
    torch.cuda.empty_cache()
    return res


def train_summary(ctx, tr):
    """The `train` extra key of the default line (and the body of the --mode train line)."""
    world, traffic = ctx.world, load_traffic()
    bwd = [r for r in tr["records"] if r[3][0] == "bwd"]
    fwd = [r for r in tr["records"] if r[3][0] == "fwd"]
    best = tr["t_graph"] or tr["t_res"]
    return {
        "metric": TRAIN_METRIC, "unit": "images/s", "steps": tr["steps"],
        "value": tr["n_img"] / best, "ms_per_step": best / tr["steps"] * 1e3,
        "value_is": "CUDA-graph replay of the whole step" if tr["t_graph"] else "eager step",
        "eager": {"value": tr["n_img"] / tr["t_res"], "ms_per_step": tr["t_res"] / tr["steps"] * 1e3},
        "e2e": {"value": tr["n_img"] / tr["t_e2e"], "unit": "images/s", "h2d_bytes_per_step": tr["h2d"],
                "d2h_bytes_per_step": tr["d2h"]},
        "graph_replay": tr["graphed"], "graph_error": tr["graph_error"], "ddp_replicas_in_sync": tr["in_sync"],
        "bf16_grad_hook": tr["bf16_grad_hook"],
        "parallelism": (f"DDP x{world} (stock DistributedDataParallel, bucketed NCCL gradient all-reduce"
                        + (", captured in the step's CUDA graph)" if tr["graphed"] else ", eager)")) if world > 1 else "single GPU",
        "workload": f"NewCRFs {ctx.args.version} training step, batch {ctx.args.batch}/GPU, 3x480x640 NYU-shape synthetic images "
                    "and depth, torch.autocast(bf16), random-init weights",
        "gpu_launches": tr["launches"], "clocks": tr["clocks"],
        "roofline_backward": with_graph_timing(
            with_bracket_probe_of(bwd, True, "swin_attn backward (all window-attention backward launches of the step)", tr["steps"], tr["t_res"], traffic),
            [r for r in tr["graph_records"] if r[3][0] == "bwd"],
            "swin_attn backward (all window-attention backward launches of the step)", tr["steps"], tr["t_graph_timed"], traffic),
        "roofline_forward": with_graph_timing(
            with_bracket_probe_of(fwd, False, "swin_attn forward (all window-attention forward launches of the step)", tr["steps"], tr["t_res"], traffic),
            [r for r in tr["graph_records"] if r[3][0] == "fwd"],
            "swin_attn forward (all window-attention forward launches of the step)", tr["steps"], tr["t_graph_timed"], traffic),
    }


def run_ours(args, env):
    ctx = Ctx(args, env)
    traffic = load_traffic()
    if args.mode == "train":
        tr = train_leg(ctx, args.steps, args.warmup)
        if ctx.rank != 0:
            return
        s = train_summary(ctx, tr)
        line = {
            "metric": TRAIN_METRIC, "value": tr["n_img"] / tr["t_res"], "unit": "images/s", "n_gpus": ctx.world,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": tr["t_res"] / args.steps * 1e3,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "bf16", "data": "synthetic",
            "config": {"workload": s["workload"], "parallelism": s["parallelism"], "l2_flush_between_steps": True,
                       "timing": "CUDA events per step, max over ranks",
                       "e2e_timing": "per step: H2D -> step -> D2H on one stream" + (
                           "; the step is replayed from a CUDA graph (harness.GraphedTrainStep), while `value` is the eager "
                           "step so that the attention launches can be bracketed by CUDA events for the roofline"
                           if tr["graphed"] else "")},
            "e2e": s["e2e"], "gpu_launches": tr["launches"], "clocks": tr["clocks"], "roofline": s["roofline_backward"],
            "roofline_forward": s["roofline_forward"],
        }
        lin = roofline_linear_of([r for r in tr["records"] if r[3][0] == "lin"], args.steps, tr["t_res"])
        if lin:
            line["roofline_linear"] = lin
        if tr["t_graph"]:
            line["graph_replay"] = {"value": tr["n_img"] / tr["t_graph"], "unit": "images/s",
                                    "ms_per_step": tr["t_graph"] / args.steps * 1e3,
                                    "note": "device-resident like `value`, the whole training step replayed from a CUDA graph "
                                            "(harness.GraphedTrainStep)", "ddp_replicas_in_sync": tr["in_sync"]}
        if tr["graph_error"]:
            line["graph_error"] = tr["graph_error"]
        emit(line)
        return

    inf = infer_leg(ctx)
    line = None
    if ctx.rank == 0:
        fwd_recs = [r for r in inf["records"] if r[3][0] == "fwd"]
        n_img, t_res = inf["n_img"], inf["t_res"]
        line = {
            "metric": METRIC, "value": n_img / t_res, "unit": "images/s", "n_gpus": ctx.world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": t_res / args.steps * 1e3, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "bf16", "data": "synthetic",
            "config": {"workload": (f"NewCRFs {args.version} full-model inference, batch {args.batch}/GPU, 3x352x1216 "
                                    "KITTI-shape synthetic images, torch.autocast(bf16), random-init weights"),
                       "parallelism": f"replicas x{ctx.world} (batch data-parallel, no data-path collective)",
                       "l2_flush_between_steps": True, "timing": "CUDA events per step, max over ranks",
                       "e2e_timing": "median of 5 regions, each one event pair around K steps of harness.infer_host_batches (pinned "
                                     "host in / out, copies of neighbouring steps overlapped on copy streams, L2 flush inside"
                                     + (")" if args.no_graph else "; the forward is replayed from a CUDA graph, "
                                        "harness.GraphedInference, while `value` is the eager forward so that the attention "
                                        "launches can be bracketed by CUDA events for the roofline; `e2e_eager` is the same loop "
                                        "with eager launches)")},
            "e2e": {"value": n_img / inf["t_e2e"], "unit": "images/s", "h2d_bytes_per_step": inf["h2d"],
                    "d2h_bytes_per_step": inf["d2h"], "regions_images_per_s": inf["e2e_regions"]},
            "gpu_launches": inf["launches"], "clocks": inf["clocks"],
            "roofline": with_graph_timing(
                with_bracket_probe_of(fwd_recs, False, "swin_attn forward (all window-attention forward launches of the step)", args.steps, t_res, traffic),
                [r for r in inf["graph_records"] if r[3][0] == "fwd"],
                "swin_attn forward (all window-attention forward launches of the step)", args.steps, inf["t_graph_timed"], traffic),
        }
        if inf["t_e2e_eager"]:
            line["e2e_eager"] = {"value": n_img / inf["t_e2e_eager"], "unit": "images/s",
                                 "note": "the e2e serving loop with the forward launched eagerly (median of 3 regions)"}
        lin = roofline_linear_of([r for r in inf["records"] if r[3][0] == "lin"], args.steps, t_res)
        if lin:
            line["roofline_linear"] = lin
        if inf["t_graph"]:
            line["graph_replay"] = {"value": n_img / inf["t_graph"], "unit": "images/s",
                                    "ms_per_step": inf["t_graph"] / args.steps * 1e3,
                                    "note": "device-resident like `value`, forward replayed from a CUDA graph (harness.GraphedInference)"}
    inf = None

    # From here on nothing may cost the line: every extra leg is guarded, and a watchdog emits what there is if the
    # training leg (NCCL collectives captured in a CUDA graph at N > 1) should ever hang.
    emitted = threading.Event()

    def emit_once():
        if not emitted.is_set():
            emitted.set()
            if ctx.rank == 0:
                emit(line)

    def watchdog():
        if ctx.rank == 0 and line is not None:
            line.setdefault("train", {"error": f"training leg did not finish within {TRAIN_LEG_TIMEOUT_S:.0f} s"})
        emit_once()
        os._exit(0)

    if ctx.world == 1 and not args.no_reference_gpu:
        try:
            line["reference_gpu_eager"] = reference_gpu_eager(args, ctx.dev, ctx.flush)
        except Exception as e:
            line["reference_gpu_eager"] = {"error": f"{type(e).__name__}: {e}"[:300]}
    if not args.no_train_leg:
        timer = threading.Timer(TRAIN_LEG_TIMEOUT_S, watchdog)
        timer.daemon = True
        timer.start()
        try:
            tr = train_leg(ctx, args.steps, max(args.warmup, 3))
            if ctx.rank == 0:
                line["train"] = train_summary(ctx, tr)
        except Exception as e:
            if ctx.rank == 0:
                line["train"] = {"error": f"{type(e).__name__}: {e}"[:300]}
            timer.cancel()
            emit_once()
            os._exit(0)   # the process group may be wedged after a failed collective: leave without the NCCL teardown
        timer.cancel()
    if ctx.rank == 0 and ctx.world == 1 and not args.no_cpu_baseline:
        try:
            line["cpu_baseline"] = cpu_reference(args.version, 5, 2, os.cpu_count() or 1)
        except Exception as e:
            line["cpu_baseline"] = {"error": f"{type(e).__name__}: {e}"[:300]}
    emit_once()


def emit(line):
    """The one JSON line, on the process's original stdout."""
    print(json.dumps(line), file=_STDOUT, flush=True)


_STDOUT = sys.stdout


def main():
    global _STDOUT
    args = parse()
    # stdout carries exactly one line: banners that libraries write to fd 1 (NCCL's version line) go to stderr
    sys.stdout.flush()
    _STDOUT = os.fdopen(os.dup(1), "w")
    os.dup2(2, 1)
    env = dist_env()
    if args.impl == "reference":
        run_reference(args, env["rank"])
        return
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device (there is no CPU fallback); use --impl reference for the CPU arm")
    if env["world"] > 1 and not args.no_graph and not args.no_graph_ddp:
        # precondition of capturing NCCL collectives in a CUDA graph (torch's DDP + whole-network-capture recipe)
        # (torchrun exports TORCH_NCCL_ASYNC_ERROR_HANDLING=1 to its workers: override, not setdefault)
        os.environ["TORCH_NCCL_ASYNC_ERROR_HANDLING"] = "0"
    from mde_b200.harness import init_distributed
    env = init_distributed("nccl")
    run_ours(args, env)
    if env["world"] > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
