"""Host-side data-parallel logic on CPU with world_size-2 gloo (no GPU, no CUDA extension compute):
rendezvous from the torchrun environment, contiguous batch sharding, metric all-reduce
(utils/dist_utils.py:45-64 semantics) and DDP gradient equivalence of ``harness.Trainer``."""
import os
import socket
import sys

import pytest
import torch
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, out):
    sys.path.insert(0, ROOT)
    os.environ.update(RANK=str(rank), LOCAL_RANK=str(rank), WORLD_SIZE=str(world), MASTER_ADDR="127.0.0.1",
                      MASTER_PORT=str(port))
    import torch.distributed as dist
    from mde_b200 import harness
    env = harness.init_distributed("gloo")
    assert env == dict(rank=rank, local_rank=rank, world=world) and dist.get_backend() == "gloo"

    # batch sharding + metric reduction
    full = torch.arange(8 * 3, dtype=torch.float32).view(8, 3)
    mine = harness.shard_batch(full, rank, world)
    assert mine.shape[0] == 4 and torch.equal(mine, full[4 * rank:4 * rank + 4])
    mean = harness.all_reduce_mean(mine.mean())
    assert torch.allclose(mean, full.mean())

    # DDP: per-rank half batches must give the full-batch gradient step (SILog per_image averages images)
    torch.manual_seed(0)
    net = torch.nn.Sequential(torch.nn.Conv2d(3, 4, 3, padding=1), torch.nn.ReLU(), torch.nn.Conv2d(4, 1, 3, padding=1),
                              torch.nn.Softplus())
    ref = torch.nn.Sequential(torch.nn.Conv2d(3, 4, 3, padding=1), torch.nn.ReLU(), torch.nn.Conv2d(4, 1, 3, padding=1),
                              torch.nn.Softplus())
    ref.load_state_dict(net.state_dict())
    images = harness.synthetic_images(4, 8, 12, seed=3)
    gt = harness.synthetic_depth(4, 8, 12, 10.0, seed=4, sparsity=0.3)
    tr = harness.Trainer(net, lr=1e-2, grad_norm=0.0, bf16=False, ddp=True)
    loss = tr.step(harness.shard_batch(images, rank, world), harness.shard_batch(gt, rank, world))
    tr_ref = harness.Trainer(ref, lr=1e-2, grad_norm=0.0, bf16=False, ddp=False)
    loss_ref = tr_ref.step(images, gt)
    assert torch.allclose(harness.all_reduce_mean(loss), loss_ref, atol=1e-6)
    for a, b in zip(net.parameters(), ref.parameters()):
        assert torch.allclose(a, b, atol=1e-6), "DDP step differs from the full-batch step"
    # the invariant DDP maintains, as bench.py checks it after replaying a captured DDP step
    assert harness.replicas_in_sync(net)
    with torch.no_grad():
        next(net.parameters()).add_(float(rank))          # rank 1 drifts
    assert not harness.replicas_in_sync(net)
    dist.barrier()
    dist.destroy_process_group()
    out.put(rank)


def test_gloo_world2_sharding_allreduce_and_ddp_step():
    ctx = mp.get_context("spawn")
    out = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, out)) for r in range(2)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(timeout=240)
    assert all(p.exitcode == 0 for p in procs), [p.exitcode for p in procs]
    assert sorted(out.get(timeout=5) for _ in procs) == [0, 1]


def test_dist_env_defaults_single_process(monkeypatch):
    sys.path.insert(0, ROOT)
    for k in ("RANK", "LOCAL_RANK", "WORLD_SIZE"):
        monkeypatch.delenv(k, raising=False)
    from mde_b200 import harness
    assert harness.dist_env() == dict(rank=0, local_rank=0, world=1)
    x = torch.ones(3)
    assert harness.all_reduce_mean(x) is x


def test_drop_path_gates_are_drawn_once_per_forward_host_logic():
    """layers.draw_drop_path_gates on CPU tensors: two rows per module and forward (a block gates its attention branch and its
    MLP branch), values 0 or 1/keep rounded to the requested dtype, modules with rate 0 untouched, a module left without
    rows draws for itself, eval mode never gates."""
    sys.path.insert(0, ROOT)
    from mde_b200 import layers
    root = torch.nn.Sequential(layers.DropPath(0.1), layers.DropPath(0.3), layers.DropPath(0.0)).train()
    torch.manual_seed(0)
    layers.draw_drop_path_gates(root, 4096, "cpu", torch.bfloat16)
    x = torch.ones(4096, 3, 2)
    assert root[2]._banked is None and root[2].gate(x) is None
    for m in (root[0], root[1]):
        want = float(torch.tensor(1.0 / (1.0 - m.drop_prob)).bfloat16())
        assert len(m._banked) == 2
        a, b = m.gate(x), m.gate(x)
        assert not m._banked and a.shape == b.shape == (4096, 1, 1) and a.dtype == torch.float32
        assert set(a.unique().tolist()) == {0.0, want} and not torch.equal(a, b)
        assert abs(float((a > 0).float().mean()) - (1.0 - m.drop_prob)) < 0.03
        c = m.gate(x)                                    # a third use in the same forward: the module's own draw
        assert c.shape == (4096, 1, 1) and abs(float((c > 0).float().mean()) - (1.0 - m.drop_prob)) < 0.03
        y = m(x)                                         # forward keeps the dtype of x
        assert y.dtype == x.dtype and set(y.unique().tolist()) <= {0.0, float(torch.tensor(1.0 / (1.0 - m.drop_prob)))}
    layers.draw_drop_path_gates(root, 8, "cpu", torch.float32)
    assert root[0]._banked[0].shape == (8,) and root[0].gate(torch.ones(4, 2)).shape == (4, 1)   # wrong batch: own draw
    root.eval()
    assert root[0].gate(x) is None and torch.equal(root[0](x), x)
