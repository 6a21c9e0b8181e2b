"""TEST / BASELINE INFRASTRUCTURE ONLY -- recipe for ``oracle/_ref/``: the reference's own Python package, made available
to the GPU box.

    python oracle/make_ref.py            # build container only (needs /root/reference)

The reference (IGNF/MAESTRO) is pure Python over eager PyTorch; ``/root/reference`` does not exist on the GPU box, so
the CPU arm of ``bench.py`` (``--impl reference`` and ``cpu_baseline``) could otherwise only time the oracle port.  This
recipe mirrors the ``maestro`` package (``*.py`` only, unmodified) from where it lies under ``/root/reference`` into
``oracle/_ref/maestro``.  ``oracle/_ref/`` is git-ignored (the sources never enter this repository's history) and NOT
gpurun-ignored, so the copy travels with the snapshot like the built ``.so`` does.  It is imported through
``oracle/ref_shim.py`` (arithmetic-free stubs for the absent third-party packages; ``vit_pytorch.vit`` served by the
restatement ``oracle/vit_restated.py``) and only by ``bench.py``'s CPU arms and the tests -- never by the product.
"""

from __future__ import annotations

import os
import shutil
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
SRC = os.environ.get("MAESTRO_REFERENCE_ROOT", "/root/reference")
DST = os.path.join(HERE, "_ref")


def make(verbose: bool = True) -> bool:
    src = os.path.join(SRC, "maestro")
    if not os.path.isdir(src):
        if verbose:
            print(f"make_ref: {src} absent (GPU box?): keeping whatever is in {DST}")
        return os.path.isdir(os.path.join(DST, "maestro"))
    dst = os.path.join(DST, "maestro")
    if os.path.isdir(dst):
        shutil.rmtree(dst)
    n = 0
    for root, dirs, files in os.walk(src):
        dirs[:] = [d for d in dirs if d != "__pycache__"]
        rel = os.path.relpath(root, src)
        os.makedirs(os.path.join(dst, rel), exist_ok=True)
        for f in files:
            if f.endswith(".py"):
                shutil.copy2(os.path.join(root, f), os.path.join(dst, rel, f))
                n += 1
    if verbose:
        print(f"make_ref: mirrored {n} reference modules into {dst} (git-ignored)")
    return True


if __name__ == "__main__":
    sys.exit(0 if make() else 1)
