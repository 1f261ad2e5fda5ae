"""Launch timing inside a CUDA-graph replay (ops.capture_graph_timing): the library records its event pair as external
event-record nodes when the stream is capturing, so bench.py can time the attention launches inside the replayed step."""
import pytest
import torch

from mde_b200 import _lib, ops
from oracle.synthetic import randn


@pytest.mark.gpu
def test_launches_captured_into_a_graph_are_timed_on_every_replay():
    dev = "cuda"
    nh, c = 4, 128
    qkv = randn("gt.qkv", 2, 28, 35, 3 * c).to(dev).bfloat16()
    table = (0.1 * randn("gt.t", 169, nh)).to(dev)
    pad = (0.1 * randn("gt.p", 3 * c)).to(dev).bfloat16()

    def fwd():
        return torch.ops.swin_attn.fwd(qkv, None, table, pad, None, 7, 3, nh, 32 ** -0.5, False, _lib.IMPL_AUTO, None, 0.0, 0)[0]

    want = fwd()
    ops.clear_graph_timing()
    side = torch.cuda.Stream()
    side.wait_stream(torch.cuda.current_stream())
    ops.capture_graph_timing(True)
    try:
        with torch.cuda.stream(side):
            fwd()                                   # eager launch in graph mode: not timed
        torch.cuda.current_stream().wait_stream(side)
        assert ops.read_graph_timing() == []
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g):
            out = fwd()
            out2 = fwd()
    finally:
        ops.capture_graph_timing(False)
    fwd()                                           # timing off again: adds no record
    for _ in range(2):                              # every replay re-records the same event pairs
        out.zero_()
        g.replay()
        torch.cuda.synchronize()
        recs = ops.read_graph_timing()
        assert len(recs) == 2
        for ms, nbytes, flops, tag in recs:
            assert 0.0 < ms < 5.0 and nbytes > 0 and flops > 0
            assert tag == ("fwd", 7, 3, nh, 28, 35)
        assert torch.equal(out, want) and torch.equal(out2, want)
    del g
    ops.clear_graph_timing()
    assert ops.read_graph_timing() == []
