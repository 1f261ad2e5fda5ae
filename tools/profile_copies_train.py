#!/usr/bin/env python
"""Which tensors does a training step copy / cast?  aten::copy_ grouped by input shapes, one NewCRFs Swin-L step (480x640, batch 8)."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402
from torch.profiler import ProfilerActivity, profile  # noqa: E402
from mde_b200.harness import Trainer, build_model, synthetic_depth, synthetic_images  # noqa: E402

dev = torch.device("cuda")
model = build_model("large07", 80.0, seed=0, device=dev)
images = synthetic_images(8, 480, 640, seed=0).to(dev)
tr = Trainer(model.train())
gt = synthetic_depth(8, 480, 640, 10.0).to(dev)
for _ in range(3):
    tr.step(images, gt)
torch.cuda.synchronize()
with profile(activities=[ProfilerActivity.CPU, ProfilerActivity.CUDA], record_shapes=True, with_stack=False) as prof:
    tr.step(images, gt)
    torch.cuda.synchronize()
rows = [e for e in prof.key_averages(group_by_input_shape=True) if e.key in ("aten::copy_", "aten::_to_copy", "aten::fill_", "aten::add_", "aten::mul", "aten::div")]
rows.sort(key=lambda e: -e.self_device_time_total)
tot = {}
for e in rows:
    tot[e.key] = tot.get(e.key, 0) + e.self_device_time_total
print({k: round(v / 1e3, 3) for k, v in tot.items()}, "ms")
for e in rows[:70]:
    print(f"{e.key:16s} n={e.count:4d} cuda={e.self_device_time_total / 1e3:8.3f} ms  {str(e.input_shapes)[:110]}")
