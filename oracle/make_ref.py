#!/usr/bin/env python
"""ORACLE — test / baseline infrastructure only (never imported by the product path).

Stages the reference's OWN hot-path sources for the GPU box.  The reference is pure Python (there is nothing to
compile), and /root/reference does not exist on the GPU box, so this recipe copies the five files SURVEY.md 8(a)
cites -- from where they lie under /root/reference, unmodified -- into the git-ignored ``oracle/_ref/`` (it travels
with ``gpurun`` like a built ``.so``; it never enters the history):

    seva/__init__.py  seva/model.py  seva/sampling.py  seva/geometry.py
    seva/modules/__init__.py  seva/modules/layers.py  seva/modules/transformer.py

``bench.py --impl reference`` and the ``gpu_library_baseline`` leg then drive the reference's own
``EulerEDMSampler.sampler_step`` -> ``DiscreteDenoiser`` -> ``SGMWrapper(Seva)`` -> ``MultiviewCFG`` (fp32 on the
host cores; bf16 autocast + its forced flash SDPA on the GPU).  ``roma`` (absent; used only by trajectory presets,
geometry.py:415) is stubbed at import time by ``oracle/refimport.py``.

    python oracle/make_ref.py            # idempotent; prints what it staged
"""
from __future__ import annotations

import hashlib
import json
import os
import shutil
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
SRC = os.environ.get("SEVA_REFERENCE_ROOT", "/root/reference")
DST = os.path.join(HERE, "_ref")
FILES = ["seva/__init__.py", "seva/model.py", "seva/sampling.py", "seva/geometry.py", "seva/modules/__init__.py",
         "seva/modules/layers.py", "seva/modules/transformer.py"]


def stage(verbose: bool = True) -> bool:
    """Copy the hot-path files; returns False (and leaves any earlier copy alone) when the reference is absent."""
    if not os.path.isdir(os.path.join(SRC, "seva")):
        if verbose:
            print(f"make_ref: {SRC} not present (GPU box?) - keeping {DST} as it is")
        return os.path.isdir(os.path.join(DST, "seva"))
    manifest = {}
    for rel in FILES:
        src, dst = os.path.join(SRC, rel), os.path.join(DST, rel)
        os.makedirs(os.path.dirname(dst), exist_ok=True)
        if os.path.exists(src):
            shutil.copyfile(src, dst)
            with open(src, "rb") as f:
                manifest[rel] = hashlib.sha256(f.read()).hexdigest()
        else:  # package markers may be absent (namespace packages)
            open(dst, "a").close()
            manifest[rel] = None
    with open(os.path.join(DST, "MANIFEST.json"), "w") as f:
        json.dump({"source": SRC, "sha256": manifest}, f, indent=1)
    if verbose:
        print(f"make_ref: staged {len(FILES)} files from {SRC} into {DST}")
    return True


if __name__ == "__main__":
    sys.exit(0 if stage() else 1)
