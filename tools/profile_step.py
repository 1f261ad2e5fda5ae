#!/usr/bin/env python
"""torch.profiler breakdown of one NewCRFs Swin-L inference (or training) step: CUDA time by op."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402
from torch.profiler import ProfilerActivity, profile  # noqa: E402
from mde_b200.harness import build_model, infer, prepare_inference, synthetic_images  # noqa: E402

train = len(sys.argv) > 1 and sys.argv[1] == "train"
dev = torch.device("cuda")
model = build_model("large07", 80.0, seed=0, device=dev)
images = synthetic_images(8, 480 if train else 352, 640 if train else 1216, seed=0).to(dev)
if train:
    from mde_b200.harness import Trainer, synthetic_depth
    tr = Trainer(model.train())
    gt = synthetic_depth(8, 480, 640, 10.0).to(dev)
    step = lambda: tr.step(images, gt)  # noqa: E731
else:
    prepare_inference(model)   # what bench.py serves: bf16 GEMM / conv weights, no per-call casts
    step = lambda: infer(model, images)  # noqa: E731
for _ in range(3):
    step()
torch.cuda.synchronize()
with profile(activities=[ProfilerActivity.CPU, ProfilerActivity.CUDA], record_shapes=True) as prof:
    step()
    torch.cuda.synchronize()
print(prof.key_averages().table(sort_by="self_cuda_time_total", row_limit=40, max_name_column_width=60))
print(prof.key_averages(group_by_input_shape=True).table(sort_by="self_cuda_time_total", row_limit=90, max_name_column_width=40,
                                                         max_shapes_column_width=70))
