"""ORACLE — test infrastructure only.  Imports the real reference from /root/reference (build container only;
the path does not exist on the GPU box) with empty stand-ins for optional UI / IO dependencies it does not
need on the hot path."""
from __future__ import annotations

import os
import sys
import types

REFERENCE_ROOT = os.environ.get("SEVA_REFERENCE_ROOT", "/root/reference")
STAGED_ROOT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "_ref")  # oracle/make_ref.py (hot path only)


def available() -> bool:
    return os.path.isdir(os.path.join(REFERENCE_ROOT, "seva"))


def staged_available() -> bool:
    return os.path.exists(os.path.join(STAGED_ROOT, "seva", "sampling.py"))


def import_staged():
    """The reference's hot-path modules from ``oracle/_ref`` (the copy ``oracle/make_ref.py`` stages for the GPU box;
    the full tree at /root/reference is preferred when it exists).  Used by bench.py's reference arm and
    ``gpu_library_baseline`` leg only."""
    root = REFERENCE_ROOT if available() else STAGED_ROOT
    if not (available() or staged_available()):
        raise RuntimeError(f"reference not found at {REFERENCE_ROOT} nor staged at {STAGED_ROOT} (run oracle/make_ref.py)")
    if "roma" not in sys.modules:
        try:
            __import__("roma")
        except Exception:
            sys.modules["roma"] = types.ModuleType("roma")
    if root not in sys.path:
        sys.path.insert(0, root)
    import seva.geometry as geometry
    import seva.model as model
    import seva.sampling as sampling

    return types.SimpleNamespace(model=model, sampling=sampling, geometry=geometry, root=root)


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


def import_eval():
    """`seva.eval` (chunk planner / scene loop bookkeeping).  Needs attribute-bearing stand-ins for colorama and
    stubs for gradio / imageio / torchvision / PIL when those are absent."""
    import_reference()
    col = sys.modules.get("colorama")
    if col is not None and not hasattr(col, "Fore"):
        col.Fore = types.SimpleNamespace(RED="", GREEN="")
        col.Style = types.SimpleNamespace(RESET_ALL="")
        col.init = lambda **k: None
    gr = sys.modules.get("gradio")
    if gr is not None and not hasattr(gr, "Progress"):
        gr.Progress = type("Progress", (), {})
    iio = sys.modules.get("imageio")
    if iio is not None and not hasattr(iio, "v3"):
        iio.v3 = sys.modules["imageio.v3"]
    for name in ("torchvision", "torchvision.transforms", "torchvision.transforms.functional", "PIL", "PIL.Image"):
        if name not in sys.modules:
            try:
                __import__(name)
            except Exception:
                sys.modules[name] = types.ModuleType(name)
                parent, _, child = name.rpartition(".")
                if parent:
                    setattr(sys.modules[parent], child, sys.modules[name])
    import seva.eval as E

    return E
