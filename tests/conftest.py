import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

GOLDEN = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


def pytest_collection_modifyitems(config, items):
    import torch
    # parity is judged against fp32 CPU outputs of the reference: keep cuDNN / cuBLAS
    # in true fp32 (PyTorch enables TF32 convolutions by default)
    torch.backends.cudnn.allow_tf32 = False
    torch.backends.cuda.matmul.allow_tf32 = False
    if torch.cuda.is_available():
        return
    skip = pytest.mark.skip(reason="no CUDA device")
    for item in items:
        if "gpu" in item.keywords:
            item.add_marker(skip)


@pytest.fixture(scope="session")
def golden():
    import torch

    def load(name):
        return torch.load(os.path.join(GOLDEN, name + ".pt"), map_location="cpu", weights_only=False)
    return load


def rel_err(a, b):
    """max |a-b| / max |b| (the max-normalised metric of SURVEY.md F5)."""
    a, b = a.detach().double().cpu(), b.detach().double().cpu()
    return float((a - b).abs().max() / b.abs().max().clamp_min(1e-30))


_ERRLOG = os.path.join(ROOT, "gpurun_out", "parity_errors.jsonl")


def check(err, bound, what):
    """assert err < bound, with the achieved error in the message AND appended to gpurun_out/parity_errors.jsonl, so
    every tolerance in the suite can be audited against what was measured (VERDICT r1 weak #1c)."""
    import json
    err = float(err)
    try:
        os.makedirs(os.path.dirname(_ERRLOG), exist_ok=True)
        test = os.environ.get("PYTEST_CURRENT_TEST", "").split(" ")[0]
        with open(_ERRLOG, "a") as f:
            f.write(json.dumps({"test": test, "what": what, "err": err, "bound": bound}) + "\n")
    except OSError:
        pass
    assert err < bound, f"{what}: measured {err:.3e} >= bound {bound:.1e}"
