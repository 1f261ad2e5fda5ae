#!/usr/bin/env python
"""Headline benchmark: NewCRFs Swin-L (``large07``) inference images/s at
352x1216, bf16 autocast, batch 8 per GPU, synthetic KITTI-shape input
(BASELINE.json configs[1]).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]

One JSON line on rank 0.  ``value`` = device-resident throughput, ``e2e`` = the
same through ``harness.infer`` with pinned-host images copied in and the depth
map copied out every step.  ``roofline`` is for the window-attention forward
kernel (all its launches inside the timed region, CUDA events on the launching
stream); ``cpu_baseline`` / ``--impl reference`` time the oracle's CPU
restatement of the reference model (the reference is Python + missing deps
and cannot travel to the GPU box).
"""
from __future__ import annotations

import argparse
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


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--mode", default="infer", choices=["infer", "train"],
                    help="infer: BASELINE configs[1] (the headline metric); train: configs[2], fwd+SILog+bwd+AdamW at 480x640")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-graph", action="store_true", help="e2e leg launches the forward eagerly instead of replaying a CUDA graph")
    ap.add_argument("--graph-ddp", action="store_true",
                    help="--mode train at N > 1: capture the DDP step (with its NCCL all-reduce) in a CUDA graph as well; opt-in "
                         "(verified at 2 GPUs; the default multi-GPU training bench stays eager)")
    ap.add_argument("--version", default=VERSION)
    ap.add_argument("--batch", type=int, default=BATCH)
    return ap.parse_args()


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

    def summary(self):
        return {"sm_mhz": statistics.median(self.samples) if self.samples else None,
                "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons)}


def cpu_model_and_sample(version):
    from oracle.cpu_model import build_newcrfs
    from mde_b200.harness import synthetic_images
    model = build_newcrfs(version, seed=0, max_depth=MAX_DEPTH)
    return model, synthetic_images(1, HEIGHT, WIDTH, seed=0)


def time_cpu(model, img, steps, warmup):
    with torch.no_grad():
        for _ in range(warmup):
            model(img)
        t0 = time.perf_counter()
        for _ in range(steps):
            model(img)
        dt = time.perf_counter() - t0
    return img.shape[0] * steps / dt, dt / steps


def run_reference(args, rank):
    """Reference arm: the reference's CPU implementation of the path (oracle port;
    see module docstring), every host thread, one image of the workload per step."""
    if rank != 0:
        return
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    model, img = cpu_model_and_sample(args.version)
    ips, spstep = time_cpu(model, img, args.steps, args.warmup)
    line = {
        "impl": "reference", "metric": METRIC, "value": ips, "unit": "images/s", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": spstep * 1e3, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": f"NewCRFs {args.version} full-model inference, 352x1216 KITTI shape",
                   "sample": "1 image per step on the host CPU"},
        "cpu_baseline": {"value": ips, "unit": "images/s", "cores": cores, "kind": "port",
                         "sample": "1 image (of the batch-8 workload) per step, fp32, torch CPU"},
        "e2e": {"value": ips, "unit": "images/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    emit(line)


def run_ours(args, env):
    from mde_b200 import _lib, ops
    from mde_b200.harness import GraphedInference, build_model, infer, infer_host_batches, prepare_inference, synthetic_images

    rank, world, local = env["rank"], env["world"], env["local_rank"]
    dev = torch.device("cuda", local)
    torch.cuda.set_device(dev)
    train = args.mode == "train"
    if train and world > 1 and args.graph_ddp and not args.no_graph:
        # the DDP step is captured in a CUDA graph (harness.GraphedTrainStep): the whole job runs on one non-default stream
        torch.cuda.set_stream(torch.cuda.Stream(dev))
    height, width, max_depth = (480, 640, 10.0) if train else (HEIGHT, WIDTH, MAX_DEPTH)
    model = build_model(args.version, max_depth, seed=0, device=dev)
    host = synthetic_images(args.batch, height, width, seed=rank, pin=True)
    images = host.to(dev)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)  # > 126 MB L2

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed_steps(step_fn, steps):
        """K steps, each bracketed by its own CUDA events with an L2 flush before it."""
        evs = []
        for _ in range(steps):
            flush.zero_()
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record()
            step_fn()
            b.record()
            evs.append((a, b))
        torch.cuda.synchronize()
        return sum(a.elapsed_time(b) for a, b in evs) / 1e3

    if train:
        from mde_b200.harness import Trainer, synthetic_depth
        model.train()
        graph_step = not args.no_graph and (world == 1 or args.graph_ddp)
        trainer = Trainer(model, ddp=world > 1, device_ids=[local] if world > 1 else None, capturable=graph_step)
        gt_host = synthetic_depth(args.batch, height, width, max_depth, seed=100 + rank).pin_memory()
        gt = gt_host.to(dev)
        out_host = torch.empty(1).pin_memory()

        def step_resident():
            trainer.step(images, gt)

        def step_e2e():
            x, y = host.to(dev, non_blocking=True), gt_host.to(dev, non_blocking=True)
            out_host.copy_(trainer.step(x, y).reshape(1), non_blocking=True)
        h2d, d2h = host.numel() * 4 + gt_host.numel() * 4, 4
    else:
        prepare_inference(model)   # eval, no grads, GEMM / conv weights stored in bf16 (what autocast reads anyway)
        out_host = torch.empty(args.batch, 1, height, width).pin_memory()

        def step_resident():
            infer(model, images)

        out_hosts = [out_host, torch.empty_like(out_host).pin_memory()]
        graphed = None if args.no_graph else GraphedInference(model, images)

        def region_e2e(steps):
            """K steps through the serving loop (harness.infer_host_batches): every step's images come from pinned host
            memory and its depth map goes back to pinned host memory inside the timed region; the copies of neighbouring
            steps overlap the compute on copy streams.  One event pair around the region, L2 flushed before each forward."""
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            torch.cuda.synchronize()
            a.record()
            infer_host_batches(model, (host for _ in range(steps)), [out_hosts[i & 1] for i in range(steps)],
                               before_step=flush.zero_, graphed=graphed)
            b.record()
            torch.cuda.synchronize()
            return a.elapsed_time(b) / 1e3

        def step_e2e():
            region_e2e(1)
        h2d, d2h = host.numel() * 4, out_host.numel() * 4

    for _ in range(args.warmup):
        step_resident()
        step_e2e()
    sampler = ClockSampler(local)
    records = []
    barrier()
    if os.environ.get("BENCH_NO_CLOCK_SAMPLER") != "1":   # diagnosis only: the JSON line then carries no clock samples
        sampler.start()
    launches0 = _lib.launch_count()
    ops.set_kernel_timing(records)
    t_res = timed_steps(step_resident, args.steps)
    ops.set_kernel_timing(None)
    launches = _lib.launch_count() - launches0
    barrier()
    gstep = None
    if train:
        if graph_step:
            # single-process training: the whole step (fwd + loss + bwd + clip + AdamW) replayed from a CUDA graph
            # (harness.GraphedTrainStep); captured after the eager region, whose launches carry the roofline events
            from mde_b200.harness import GraphedTrainStep
            gstep = GraphedTrainStep(trainer, images, gt, warmup=2)

            def step_e2e():   # noqa: F811  (the end-to-end step through the graph: H2D -> static inputs -> replay -> D2H)
                x, y = host.to(dev, non_blocking=True), gt_host.to(dev, non_blocking=True)
                out_host.copy_(gstep(x, y).reshape(1), non_blocking=True)
            for _ in range(2):
                step_e2e()
        t_e2e = timed_steps(step_e2e, args.steps)
    else:
        # One event pair brackets the whole K-step serving region, so a single host hiccup (GC, the clock sampler taking
        # the GIL, a page fault in a pinned buffer) while the ~100 ms region is being enqueued lands in the number in full
        # (observed: 464 / 649 images/s on runs whose other regions gave 770-800; on a 2-GPU run regions of 1180 / 962 /
        # 318 images/s beside 1540-1600 on the next four runs of the same command).  The region is therefore timed five
        # times, each with exactly K steps and a synchronize on both sides, and the median is reported.
        e2e_regions = sorted(region_e2e(args.steps) for _ in range(5))
        t_e2e = e2e_regions[2]
    t_graph = None
    if not train and graphed is not None:
        barrier()
        t_graph = timed_steps(lambda: graphed(images), args.steps)   # device-resident, the forward replayed as one graph
    in_sync = None
    if gstep is not None:
        t_graph = timed_steps(lambda: gstep(images, gt), args.steps)  # device-resident, the training step as one graph
        if world > 1:
            # every rank trains on its own images: the replicas stay identical only if the captured step really
            # all-reduces the gradients
            from mde_b200.harness import replicas_in_sync
            in_sync = replicas_in_sync(model)
    barrier()
    sampler.stop_flag = True
    if sampler.is_alive():
        sampler.join(timeout=2)

    times = torch.tensor([t_res, t_e2e, t_graph or 0.0], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(times, op=dist.ReduceOp.MAX)
    t_res, t_e2e, t_graph = times.tolist()
    if rank != 0:
        return
    hbm, tf, which = peaks()

    def roof(recs, name):
        k_ms = sum(r[0] for r in recs)
        if not recs or not k_ms:
            return None
        k_bytes, k_flops = sum(r[1] for r in recs), sum(r[2] for r in recs)
        achieved = k_bytes / (k_ms * 1e-3) / 1e9
        return {"bound": "hbm", "kernel": name, "achieved": achieved, "peak": hbm, "unit": "GB/s", "frac": achieved / hbm,
                "peak_source": f"{which} (MEASURED_PEAKS.json hbm_gbs)",
                "traffic": (traffic.get(name.split()[1]) or {}).get("dram_bytes_per_launch"),
                "traffic_of": traffic.get(name.split()[1]),
                "launches": len(recs), "kernel_ms_per_step": k_ms / args.steps, "share_of_step": k_ms / 1e3 / t_res,
                "useful_tflops": k_flops / (k_ms * 1e-3) / 1e12,
                "tensor_frac_of_sustained_bf16": k_flops / (k_ms * 1e-3) / 1e12 / tf}
    traffic = {}
    tpath = os.path.join(ROOT, "profiles", "roofline_traffic.json")
    if os.path.exists(tpath):
        traffic = json.load(open(tpath))
    fwd_recs = [r for r in records if r[3][0] == "fwd"]
    bwd_recs = [r for r in records if r[3][0] == "bwd"]
    roof_fwd = roof(fwd_recs, "swin_attn forward (all window-attention forward launches of the step)")
    roof_bwd = roof(bwd_recs, "swin_attn backward (all window-attention backward launches of the step)")
    n_img = args.batch * world * args.steps
    if train:
        metric = "NewCRFs Swin-L training images/sec @480x640 (bf16 autocast fwd + SILog + bwd + clip + AdamW, batch 8/GPU)"
        workload = (f"NewCRFs {args.version} training step, batch {args.batch}/GPU, 3x480x640 NYU-shape synthetic images and "
                    "depth, torch.autocast(bf16), random-init weights")
        par = f"DDP x{world} (stock DistributedDataParallel, NCCL gradient all-reduce)" if world > 1 else "single GPU"
    else:
        metric = METRIC
        workload = (f"NewCRFs {args.version} full-model inference, batch {args.batch}/GPU, 3x352x1216 "
                    "KITTI-shape synthetic images, torch.autocast(bf16), random-init weights")
        par = f"replicas x{world} (batch data-parallel, no data-path collective)"
    line = {
        "metric": metric, "value": n_img / t_res, "unit": "images/s", "n_gpus": world, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": t_res / args.steps * 1e3, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "bf16", "data": "synthetic",
        "config": {"workload": workload, "parallelism": par,
                   "l2_flush_between_steps": True, "timing": "CUDA events per step, max over ranks",
                   "e2e_timing": (("per step: H2D -> step -> D2H on one stream" + ("" if gstep is None else
                                   "; the step is replayed from a CUDA graph (harness.GraphedTrainStep), while `value` is the eager "
                                   "step so that the attention launches can be bracketed by CUDA events for the roofline"))
                                  if train else
                                  "median of 5 regions, each one event pair around K steps of harness.infer_host_batches (pinned host in / out, "
                                  "copies of neighbouring steps overlapped on copy streams, L2 flush inside"
                                  + (")" if args.no_graph else "; the forward is replayed from a CUDA graph, "
                                     "harness.GraphedInference, while `value` is the eager forward so that the attention "
                                     "launches can be bracketed by CUDA events for the roofline)"))},
        "e2e": {"value": n_img / t_e2e, "unit": "images/s", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                **({} if train else {"regions_images_per_s": [n_img / t for t in e2e_regions]})},
        "gpu_launches": int(launches),
        "clocks": sampler.summary(),
        "roofline": roof_bwd if train else roof_fwd,
    }
    lin_recs = [r for r in records if r[3][0] == "lin"]
    if lin_recs:
        l_ms, l_bytes, l_flops = (sum(r[i] for r in lin_recs) for i in range(3))
        line["roofline_linear"] = {
            "bound": "tensor", "kernel": "tc_linear_kernel (fc1 + bias + GELU of every block MLP, tcgen05 GEMM with fused epilogue; CTA pairs / cta_group::2 from K = 192 up)",
            "achieved": l_flops / (l_ms * 1e-3) / 1e12, "peak": tf, "unit": "TFLOP/s", "frac": l_flops / (l_ms * 1e-3) / 1e12 / tf,
            "peak_source": f"{which} (MEASURED_PEAKS.json dense bf16)", "traffic": None,
            "algorithmic_GBps": l_bytes / (l_ms * 1e-3) / 1e9, "hbm_frac": l_bytes / (l_ms * 1e-3) / 1e9 / hbm,
            "launches": len(lin_recs), "kernel_ms_per_step": l_ms / args.steps, "share_of_step": l_ms / 1e3 / t_res,
            "by_shape": {}}
        shapes = {}
        for r in lin_recs:                                   # tag = ("lin", in_features, activation, out_features, rows, 0)
            shapes.setdefault((r[3][4], r[3][1], r[3][3]), []).append(r)
        for (rows, k, n), rs in sorted(shapes.items(), reverse=True):
            ms = sum(r[0] for r in rs)
            line["roofline_linear"]["by_shape"][f"{rows}x{k}x{n}"] = {
                "launches": len(rs), "us_per_launch": ms / len(rs) * 1e3, "TFLOPs": sum(r[2] for r in rs) / (ms * 1e-3) / 1e12,
                "GBps": sum(r[1] for r in rs) / (ms * 1e-3) / 1e9}
    if t_graph:
        line["graph_replay"] = {"value": n_img / t_graph, "unit": "images/s", "ms_per_step": t_graph / args.steps * 1e3,
                                "note": ("device-resident like `value`, the whole training step replayed from a CUDA graph "
                                         "(harness.GraphedTrainStep)" if train else
                                         "device-resident like `value`, forward replayed from a CUDA graph (harness.GraphedInference)")}
    if train:
        line["roofline_forward"] = roof_fwd
        if in_sync is not None:
            line["graph_replay"]["ddp_replicas_in_sync"] = in_sync
    if world == 1 and not args.no_cpu_baseline and not train:
        cores = os.cpu_count() or 1
        torch.set_num_threads(cores)
        cm, img = cpu_model_and_sample(args.version)
        ips, _ = time_cpu(cm, img, 1, 1)
        line["cpu_baseline"] = {"value": ips, "unit": "images/s", "cores": cores, "kind": "port",
                                "sample": "1 image of the batch-8 workload, 1 warm-up + 1 timed forward, fp32 torch CPU "
                                          "(oracle restatement of the reference model)"}
    emit(line)


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
    from mde_b200.harness import dist_env, init_distributed
    env = dist_env()
    if args.impl == "reference":
        run_reference(args, env["rank"])
        return
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device (there is no CPU fallback); use --impl reference for the CPU arm")
    if args.mode == "train" and args.graph_ddp and not args.no_graph:
        # precondition of capturing NCCL collectives in a CUDA graph (torch's DDP + whole-network-capture recipe)
        # (torchrun exports TORCH_NCCL_ASYNC_ERROR_HANDLING=1 to its workers: override, not setdefault)
        os.environ["TORCH_NCCL_ASYNC_ERROR_HANDLING"] = "0"
    env = init_distributed("nccl")
    run_ours(args, env)
    if env["world"] > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
