#!/usr/bin/env python
"""Where do the elementwise / copy kernels of one inference step come from?  torch.profiler with Python stacks,
grouped by call site, for the ops outside the attention path (copy_, add, gelu, native_layer_norm, cat)."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402
from torch.profiler import ProfilerActivity, profile  # noqa: E402
from mde_b200.harness import build_model, infer, prepare_inference, synthetic_images  # noqa: E402

dev = torch.device("cuda")
train = len(sys.argv) > 1 and sys.argv[1] == "train"
if train:
    from mde_b200.harness import Trainer, synthetic_depth
    model = build_model("large07", 10.0, seed=0, device=dev).train()
    images = synthetic_images(8, 480, 640, seed=0).to(dev)
    gt = synthetic_depth(8, 480, 640, 10.0).to(dev)
    tr = Trainer(model)
    step = lambda: tr.step(images, gt)  # noqa: E731
else:
    model = prepare_inference(build_model("large07", 80.0, seed=0, device=dev))
    images = synthetic_images(8, 352, 1216, seed=0).to(dev)
    step = lambda: infer(model, images)  # noqa: E731
for _ in range(3):
    step()
torch.cuda.synchronize()
a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
a.record()
for _ in range(3):
    step()
b.record()
torch.cuda.synchronize()
print(f"step (events, no profiler): {a.elapsed_time(b) / 3:.3f} ms")
with profile(activities=[ProfilerActivity.CPU, ProfilerActivity.CUDA], with_stack=True, record_shapes=True) as prof:
    step()
    torch.cuda.synchronize()
tot = sum(e.self_device_time_total for e in prof.key_averages() if not e.key.startswith(("aten::", "swin_attn::", "Optimizer", "Generated", "autograd::", "torch")) and "Backward" not in e.key)
print(f"kernel time total {tot / 1e3:.3f} ms")
for e in sorted(prof.key_averages(), key=lambda e: -e.self_device_time_total)[:45]:
    print(f"  {e.key[:80]:80s} {e.self_device_time_total / 1e3:8.3f} ms x{e.count}")
want = ("aten::copy_", "aten::add", "aten::add_", "aten::gelu", "aten::native_layer_norm", "aten::cat", "aten::mul", "aten::_to_copy", "aten::pixel_shuffle", "aten::cudnn_convolution", "aten::addmm", "aten::mm",
        "aten::native_layer_norm_backward", "aten::gelu_backward", "aten::sum", "aten::fill_", "aten::zero_", "aten::native_dropout", "aten::div", "aten::div_")
rows = [e for e in prof.key_averages(group_by_stack_n=12, group_by_input_shape=True) if e.key in want]
rows.sort(key=lambda e: -e.self_device_time_total)
for e in rows[:40]:
    stack = [s for s in e.stack if "mde_b200" in s or "monocular" in s or "harness" in s][:3]
    print(f"{e.key:26s} {e.self_device_time_total / 1e3:8.3f} ms  x{e.count:<4d} {str(e.input_shapes)[:70]:70s} | " + " <- ".join(s.split('/')[-1] for s in stack))
