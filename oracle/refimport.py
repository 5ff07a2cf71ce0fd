"""ORACLE — test infrastructure only.  Imports the real reference from /root/reference (build container only;
the path does not exist on the GPU box) with empty stand-ins for optional UI / IO dependencies it does not
need on the hot path."""
from __future__ import annotations

import os
import sys
import types

REFERENCE_ROOT = os.environ.get("SEVA_REFERENCE_ROOT", "/root/reference")


def available() -> bool:
    return os.path.isdir(os.path.join(REFERENCE_ROOT, "seva"))


def import_reference():
    if not available():
        raise RuntimeError(f"reference not found at {REFERENCE_ROOT}")
    for name in ("roma", "gradio", "colorama", "imageio", "imageio.v3"):
        if name not in sys.modules:
            try:
                __import__(name)
            except Exception:
                sys.modules[name] = types.ModuleType(name)
    if REFERENCE_ROOT not in sys.path:
        sys.path.insert(0, REFERENCE_ROOT)
    import seva.geometry as geometry
    import seva.model as model
    import seva.sampling as sampling

    return types.SimpleNamespace(model=model, sampling=sampling, geometry=geometry)
