"""Stage the UNMODIFIED reference Python sources into the git-ignored ``oracle/_ref/``.

TEST INFRASTRUCTURE ONLY (see oracle/__init__.py).  The reference is pure Python: its "build" is a copy.  The GPU
box has no ``/root/reference`` but receives everything under the repo that is not gpurun-ignored, so staging
``model/**.py`` and ``utils/*.py`` here (from ``__graft_entry__.build()``, in the build container) is what lets the
box run the reference's own modules: the CPU arm of ``bench.py``, the ``reference_gpu_eager`` bar and the live
parity tests import them from ``oracle/_ref`` through ``oracle/reference_shim.py``.  Nothing is edited; a manifest
with the sha256 of every staged file is written next to them (``oracle/_ref/MANIFEST.json``) so a reader can check
the copies against the upstream tree.  ``oracle/_ref/`` is listed in ``.gitignore``: reference sources never enter
the history.

    python -m oracle.stage_reference            # copy (idempotent)
"""
from __future__ import annotations

import hashlib
import json
import os
import shutil

HERE = os.path.dirname(os.path.abspath(__file__))
DEST = os.path.join(HERE, "_ref")
SOURCE = os.environ.get("MDE_REFERENCE_SOURCE", "/root/reference")
TREES = ("model", "utils")


def staged() -> bool:
    return os.path.isfile(os.path.join(DEST, "model", "NewCRFs", "NewCRFDepth.py"))


def stage(source: str = SOURCE, dest: str = DEST) -> int:
    """Copy ``source/{model,utils}/**.py`` to ``dest`` byte for byte.  Returns the number of files (0 when the
    reference tree is absent, e.g. on the GPU box, where the prebuilt copy is used as it is)."""
    if not os.path.isdir(os.path.join(source, "model", "NewCRFs")):
        return 0
    manifest = {}
    for tree in TREES:
        for root, _dirs, files in os.walk(os.path.join(source, tree)):
            for name in sorted(files):
                if not name.endswith(".py"):
                    continue
                src = os.path.join(root, name)
                rel = os.path.relpath(src, source)
                dst = os.path.join(dest, rel)
                os.makedirs(os.path.dirname(dst), exist_ok=True)
                with open(src, "rb") as f:
                    data = f.read()
                manifest[rel] = hashlib.sha256(data).hexdigest()
                if not os.path.exists(dst) or open(dst, "rb").read() != data:
                    shutil.copyfile(src, dst)
    with open(os.path.join(dest, "MANIFEST.json"), "w") as f:
        json.dump({"source": source, "files": manifest}, f, indent=1, sort_keys=True)
    return len(manifest)


if __name__ == "__main__":
    n = stage()
    print(f"staged {n} reference files into {DEST}" if n else f"no reference tree at {SOURCE}; nothing staged")
