"""Synthetic-batch inference / training harness for NewCRFs.

The reference's own entry script (``run.py``) and its loss module are absent
from the snapshot; this is NEW code reconstructed from the
configuration schema (json/kitti/newcrfs/newcrfs_github_eval.json: loss
alpha=10, beta=0.15, per_image; AdamW lr 2e-5, wd 0; grad_norm 0.1) and from
utils/dist_utils.py.  Data parallelism is stock ``DistributedDataParallel``
over NCCL (one process per GPU); the attention itself is never sharded.
"""
from __future__ import annotations

import os
from typing import Callable, Dict, Iterable, List, Optional

import weakref

import torch
import torch.distributed as dist
import torch.nn as nn

from .newcrfs import NewCRFDepth

KITTI = dict(height=352, width=1216, max_depth=80.0)
NYU = dict(height=480, width=640, max_depth=10.0)


def synthetic_images(batch: int, height: int, width: int, seed: int = 0, device="cpu", pin: bool = False):
    """ImageNet-normalised-like images (dataset/depth_dataset.py:290): N(0,1)."""
    g = torch.Generator().manual_seed(seed)
    x = torch.randn(batch, 3, height, width, generator=g)
    if pin and torch.cuda.is_available():
        x = x.pin_memory()
    return x.to(device) if device != "cpu" else x


def synthetic_depth(batch: int, height: int, width: int, max_depth: float, seed: int = 1, sparsity: float = 0.7):
    g = torch.Generator().manual_seed(seed)
    gt = 1e-3 + (max_depth - 1e-3) * torch.rand(batch, 1, height, width, generator=g)
    return gt * (torch.rand(batch, 1, height, width, generator=g) > sparsity)


class SILogLoss(nn.Module):
    """Scale-invariant log loss: alpha * sqrt(mean(d^2) - beta * mean(d)^2), d = log(pred) - log(gt),
    over valid pixels; ``per_image`` averages the per-image losses."""

    def __init__(self, alpha: float = 10.0, beta: float = 0.15, per_image: bool = True, min_depth: float = 1e-3):
        super().__init__()
        self.alpha, self.beta, self.per_image, self.min_depth = alpha, beta, per_image, min_depth

    def forward(self, pred, gt):
        valid = gt > self.min_depth
        d = torch.where(valid, torch.log(pred.float().clamp_min(self.min_depth)) - torch.log(gt.clamp_min(self.min_depth)),
                        torch.zeros_like(gt))
        dims = (1, 2, 3) if self.per_image else (0, 1, 2, 3)
        n = valid.sum(dim=dims).clamp_min(1).float()
        m1 = d.sum(dim=dims) / n
        m2 = (d * d).sum(dim=dims) / n
        return (self.alpha * torch.sqrt((m2 - self.beta * m1 * m1).clamp_min(1e-12))).mean()


def dist_env() -> Dict[str, int]:
    return dict(rank=int(os.environ.get("RANK", 0)), local_rank=int(os.environ.get("LOCAL_RANK", 0)),
                world=int(os.environ.get("WORLD_SIZE", 1)))


def init_distributed(backend: Optional[str] = None) -> Dict[str, int]:
    env = dist_env()
    if env["world"] > 1 and not dist.is_initialized():
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        os.environ.setdefault("MASTER_PORT", "29500")
        backend = backend or ("nccl" if torch.cuda.is_available() else "gloo")
        kwargs = {}
        if backend == "nccl":
            torch.cuda.set_device(env["local_rank"])
            kwargs["device_id"] = torch.device("cuda", env["local_rank"])
        dist.init_process_group(backend=backend, rank=env["rank"], world_size=env["world"], **kwargs)
    return env


def replicas_in_sync(model: nn.Module, rtol: float = 0.0) -> bool:
    """True when every rank holds the same parameters (the invariant DDP maintains: identical initial values, identical
    all-reduced gradients, identical optimizer steps): compares an fp64 checksum of every parameter across ranks."""
    if not (dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1):
        return True
    with torch.no_grad():
        chk = torch.stack([p.detach().double().sum() for p in model.parameters()])
        lo, hi = chk.clone(), chk.clone()
        dist.all_reduce(lo, op=dist.ReduceOp.MIN)
        dist.all_reduce(hi, op=dist.ReduceOp.MAX)
        return bool(((hi - lo).abs() <= rtol * hi.abs()).all())


def shard_batch(t: torch.Tensor, rank: int, world: int) -> torch.Tensor:
    """Contiguous split of the global batch, as DistributedSampler-without-shuffle would."""
    per = t.shape[0] // world
    return t[rank * per:(rank + 1) * per]


def all_reduce_mean(value: torch.Tensor) -> torch.Tensor:
    """utils/dist_utils.py:45-64 (all_reduce_tensor, SUM then / world)."""
    if dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1:
        value = value.clone()
        dist.all_reduce(value, op=dist.ReduceOp.SUM)
        value = value / dist.get_world_size()
    return value


def build_model(version: str = "large07", max_depth: float = 80.0, seed: int = 0, device="cuda") -> NewCRFDepth:
    torch.manual_seed(seed)
    return NewCRFDepth(version=version, max_depth=max_depth).to(device)


def prepare_inference(model: nn.Module, bf16: bool = True) -> nn.Module:
    """Deployment transform for ``infer``: eval mode, gradients off, and the Linear / Conv2d parameters stored in
    bf16.  Under ``torch.autocast`` those are exactly the values every GEMM / conv already reads (autocast rounds
    the fp32 parameters to bf16 on each call and drops the copies when the context exits), so outputs are bit
    identical while ~400 cast kernels per forward disappear.  Norm parameters and the relative-position bias
    tables stay fp32.  In place; returns the model."""
    model.eval()
    for p in model.parameters():
        p.requires_grad_(False)
    if bf16:
        for m in model.modules():
            if isinstance(m, (nn.Linear, nn.Conv2d)):
                m.to(torch.bfloat16)
    # (Storing the conv weights channels-last as well saves the per-call weight transposition -- 771 vs 753 images/s
    # eager -- but cuDNN then picks other kernels and the outputs are no longer bit-identical to autocast: not done.)
    return model


@torch.no_grad()
def infer(model: nn.Module, images: torch.Tensor, bf16: bool = True) -> torch.Tensor:
    """The call a user makes for inference: images (B,3,H,W) on the model's device -> depth (B,1,H,W)."""
    with torch.autocast("cuda", dtype=torch.bfloat16, enabled=bf16):
        return model(images)


class GraphedInference:
    """``infer(model, x)`` for one input shape as a CUDA graph: the ~440 launches of a forward are captured once (after
    ``prepare_inference``) and replayed as one graph launch, which removes the launch gaps between the small kernels of
    stages 2 / 3.  The forward has no host synchronisation and no data-dependent control flow, and the tensor maps of
    the tcgen05 kernels are kernel parameters encoded at capture time over the graph's private memory pool, so a replay
    computes exactly what the eager forward computes (tests/test_parity_gpu.py).  ``__call__`` copies ``x`` into the
    static input and returns the static output, valid until the next call."""

    def __init__(self, model: nn.Module, example: torch.Tensor, bf16: bool = True, warmup: int = 2):
        self.model, self.bf16 = model, bf16
        self.static_in = example.detach().clone()
        side = torch.cuda.Stream(example.device)
        side.wait_stream(torch.cuda.current_stream(example.device))
        with torch.cuda.stream(side):                      # allocator and cuBLAS / cuDNN workspaces warm before the capture
            for _ in range(warmup):
                infer(model, self.static_in, bf16)
        torch.cuda.current_stream(example.device).wait_stream(side)
        self.graph = torch.cuda.CUDAGraph()
        # thread-local capture mode: CUDA calls of other threads (the NCCL watchdog of a multi-GPU job polls events)
        # must not invalidate the capture
        with torch.cuda.graph(self.graph, capture_error_mode="thread_local"):
            self.static_out = infer(model, self.static_in, bf16)

    @torch.no_grad()
    def __call__(self, x: torch.Tensor) -> torch.Tensor:
        if x.shape != self.static_in.shape or x.dtype != self.static_in.dtype:
            raise ValueError(f"graph captured for {tuple(self.static_in.shape)} {self.static_in.dtype}, "
                             f"got {tuple(x.shape)} {x.dtype}")
        self.static_in.copy_(x)
        self.graph.replay()
        return self.static_out


class _ServingState:
    """Streams, device staging buffers and events of ``infer_host_batches`` for one (device, batch shape): created once
    and reused, so that a steady-state serving loop allocates nothing -- no caching-allocator traffic across streams (a
    ``record_stream`` block that cannot be reused yet turns into a ``cudaMalloc``, which synchronises the device), no
    stream or event creation per step."""

    N_IN, N_OUT = 3, 2

    def __init__(self, dev, x_shape, x_dtype):
        self.key = (dev, tuple(x_shape), x_dtype)
        self.up, self.down = torch.cuda.Stream(dev), torch.cuda.Stream(dev)
        self.xin = [torch.empty(x_shape, dtype=x_dtype, device=dev) for _ in range(self.N_IN)]
        self.in_ready = [torch.cuda.Event() for _ in range(self.N_IN)]   # upload of slot s finished (recorded on `up`)
        self.in_free = [torch.cuda.Event() for _ in range(self.N_IN)]    # the forward that read slot s finished (on `cur`)
        self.ystage: List[Optional[torch.Tensor]] = [None] * self.N_OUT  # fp32 device copies of the results
        self.out_done = [torch.cuda.Event() for _ in range(self.N_OUT)]  # result staged (on `cur`)
        self.out_free = [torch.cuda.Event() for _ in range(self.N_OUT)]  # its download finished (on `down`)


_SERVING: "weakref.WeakKeyDictionary[nn.Module, _ServingState]" = weakref.WeakKeyDictionary()   # not on the module: it must stay deep-copyable


@torch.no_grad()
def infer_host_batches(model: nn.Module, batches: Iterable[torch.Tensor], outs: Optional[List[torch.Tensor]] = None,
                       bf16: bool = True, before_step: Optional[Callable[[], None]] = None,
                       graphed: Optional[GraphedInference] = None) -> List[torch.Tensor]:
    """Inference over a sequence of HOST batches (pinned (B,3,H,W) tensors of one shape) -> host depth maps.

    The serving loop of a deployment: the host-to-device copy of batches i + 1 and i + 2 and the device-to-host copy of
    result i - 1 run on two copy streams while batch i computes on the current stream.  Inputs land in three rotating
    device buffers, results leave through two rotating fp32 device buffers; buffers, streams and events are created on the
    first call for a batch shape and reused afterwards (kept per model), so the steady state allocates nothing.
    ``outs`` are optional pinned result buffers, one per batch (reused by the caller); ``before_step`` is called on the
    compute stream before every forward (bench.py flushes L2 there).  ``graphed`` (a ``GraphedInference`` of ``model``
    for this batch shape) replays the captured forward instead of launching it kernel by kernel.  The last result is
    complete once the current stream has been synchronised (the function joins the download stream into it before
    returning)."""
    dev = next(model.parameters()).device
    cur = torch.cuda.current_stream(dev)
    it = iter(batches)
    first = next(it, None)
    if first is None:
        return []
    st: Optional[_ServingState] = _SERVING.get(model)
    if st is None or st.key != (dev, tuple(first.shape), first.dtype):
        st = _ServingState(dev, first.shape, first.dtype)
        _SERVING[model] = st
    st.up.wait_stream(cur)      # whatever produced / still reads the buffers on this stream is ordered before the copies
    st.down.wait_stream(cur)
    pending = [first]

    def upload(i: int) -> bool:
        """Host -> device copy of batch i into slot i % N_IN, once the forward of batch i - N_IN has read that slot."""
        b = pending.pop() if pending else next(it, None)
        if b is None:
            return False
        s = i % st.N_IN
        st.up.wait_event(st.in_free[s])
        with torch.cuda.stream(st.up):
            st.xin[s].copy_(b, non_blocking=True)
            st.in_ready[s].record(st.up)
        return True

    results: List[torch.Tensor] = []
    n_up = 0
    while n_up < st.N_IN - 1 and upload(n_up):      # two uploads in flight before the first forward
        n_up += 1
    i = 0
    while i < n_up:
        s, o = i % st.N_IN, i % st.N_OUT
        cur.wait_event(st.in_ready[s])
        if before_step is not None:
            before_step()
        y = infer(model, st.xin[s], bf16) if graphed is None else graphed(st.xin[s])
        st.in_free[s].record(cur)
        if upload(n_up):                            # slot (i + 2) % 3: free since the forward of batch i - 1
            n_up += 1
        if st.ystage[o] is None or st.ystage[o].shape != y.shape:
            st.ystage[o] = torch.empty(y.shape, dtype=torch.float32, device=dev)
        cur.wait_event(st.out_free[o])              # the download of result i - 2 has left this buffer
        st.ystage[o].copy_(y)                       # (also detaches the result from a graph's static output)
        st.out_done[o].record(cur)
        out = outs[i] if outs is not None else torch.empty(y.shape, dtype=torch.float32).pin_memory()
        st.down.wait_event(st.out_done[o])
        with torch.cuda.stream(st.down):
            out.copy_(st.ystage[o], non_blocking=True)
            st.out_free[o].record(st.down)
        results.append(out)
        i += 1
    cur.wait_stream(st.down)
    return results


class Trainer:
    """fwd + SILog + bwd (+ DDP all-reduce) + clip + AdamW, per the reference's JSON schema."""

    def __init__(self, model: nn.Module, lr: float = 2e-5, grad_norm: float = 0.1, bf16: bool = True,
                 ddp: bool = False, device_ids=None, capturable: bool = False):
        self.raw = model
        # gradients live inside the all-reduce buckets: no per-parameter copy kernel (~600 launches per step)
        if ddp and capturable and next(model.parameters()).is_cuda and \
                torch.cuda.current_stream() == torch.cuda.default_stream():
            # torch's recipe for capturing a DistributedDataParallel step builds the wrapper off the default stream; built
            # on a side stream while the steps run on the default one, every gradient accumulation pays a cross-stream
            # synchronisation in eager mode (42.3 against 39.4 ms at 2 GPUs) -- so the whole job moves to one stream
            raise ValueError("a capturable DDP Trainer must be built, and stepped, on a non-default CUDA stream: call "
                             "torch.cuda.set_stream(torch.cuda.Stream()) first")
        self.model = nn.parallel.DistributedDataParallel(model, device_ids=device_ids,
                                                         gradient_as_bucket_view=True) if ddp else model
        # one multi-tensor kernel per step; ``capturable``: the step counter lives on the device (GraphedTrainStep)
        self.opt = torch.optim.AdamW(model.parameters(), lr=lr, weight_decay=0.0, fused=next(model.parameters()).is_cuda,
                                     capturable=capturable and next(model.parameters()).is_cuda)
        self.loss_fn = SILogLoss()
        self.grad_norm, self.bf16 = grad_norm, bf16

    def step(self, images: torch.Tensor, gt: torch.Tensor) -> torch.Tensor:
        self.opt.zero_grad(set_to_none=True)
        with torch.autocast(images.device.type, dtype=torch.bfloat16, enabled=self.bf16):
            pred = self.model(images)
        loss = self.loss_fn(pred, gt)
        loss.backward()
        if self.grad_norm:
            if self.opt.defaults.get("fused"):
                # clip_grad_norm_ without its pass over the gradients: the fused AdamW kernel divides every gradient by
                # `grad_scale` on the fly (the hook torch's GradScaler uses), so the clip coefficient goes there
                grads = [p.grad for p in self.raw.parameters() if p.grad is not None]
                total = torch.linalg.vector_norm(torch.stack(torch._foreach_norm(grads)))
                self.opt.grad_scale = ((total + 1e-6) / self.grad_norm).clamp(min=1.0).float()   # 1 / clip coefficient
            else:
                torch.nn.utils.clip_grad_norm_(self.raw.parameters(), self.grad_norm)
        self.opt.step()
        return loss.detach()


class GraphedTrainStep:
    """``trainer.step(images, gt)`` for one batch shape as a CUDA graph: forward, SILog loss, backward, gradient clipping
    and the fused AdamW update (~4 000 launches, no host synchronisation, no data-dependent control flow; the tensor
    maps of the tcgen05 kernels are kernel parameters encoded at capture time over the graph's private memory pool)
    captured once and replayed as one graph launch.  The eager training step is close to launch-bound -- its time follows
    the host CPU from box to box -- and the replay is not.  With DistributedDataParallel the bucketed NCCL all-reduce is
    captured with the step (see __init__ for the preconditions).  The trainer needs ``capturable=True``.  The ``warmup`` eager steps before the capture are real training steps.
    ``__call__`` copies the batch into the static inputs and returns the static loss, valid until the next call."""

    def __init__(self, trainer: "Trainer", images: torch.Tensor, gt: torch.Tensor, warmup: int = 3):
        if isinstance(trainer.model, nn.parallel.DistributedDataParallel):
            # DDP: the bucketed NCCL all-reduce is captured with the step.  torch's recipe: async error handling of the
            # NCCL process group off before it is created, the wrapper built off the default stream (a Trainer with
            # capturable=True insists on it), and at least 11 eager DDP iterations before the capture (bucket rebuild,
            # reducer state).
            if os.environ.get("TORCH_NCCL_ASYNC_ERROR_HANDLING", os.environ.get("NCCL_ASYNC_ERROR_HANDLING")) != "0":
                raise ValueError("capturing a DDP step needs TORCH_NCCL_ASYNC_ERROR_HANDLING=0 set before init_process_group")
            warmup = max(warmup, 11)
        if not trainer.opt.defaults.get("capturable", False):
            raise ValueError("build the Trainer with capturable=True (device-side AdamW step counter)")
        if any(getattr(getattr(m, "attn_drop", None), "p", 0.0) > 0.0 for m in trainer.raw.modules()):
            # the attention-dropout seed is a host scalar drawn per call (ops.window_attention): a captured graph would
            # replay ONE mask for ever
            raise ValueError("GraphedTrainStep does not support attn_drop > 0 (the dropout seed would be baked into the graph)")
        self.trainer = trainer
        self.static_images, self.static_gt = images.detach().clone(), gt.detach().clone()
        side = torch.cuda.Stream(images.device)
        side.wait_stream(torch.cuda.current_stream(images.device))
        with torch.cuda.stream(side):               # allocator, cuBLAS / cuDNN workspaces and optimizer state warm
            for _ in range(warmup):
                trainer.step(self.static_images, self.static_gt)
        torch.cuda.current_stream(images.device).wait_stream(side)
        self.graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(self.graph, capture_error_mode="thread_local"):
            self.static_loss = trainer.step(self.static_images, self.static_gt)

    def __call__(self, images: torch.Tensor, gt: torch.Tensor) -> torch.Tensor:
        if images.shape != self.static_images.shape or gt.shape != self.static_gt.shape:
            raise ValueError(f"graph captured for {tuple(self.static_images.shape)} / {tuple(self.static_gt.shape)}, "
                             f"got {tuple(images.shape)} / {tuple(gt.shape)}")
        self.static_images.copy_(images, non_blocking=True)
        self.static_gt.copy_(gt, non_blocking=True)
        self.graph.replay()
        return self.static_loss

