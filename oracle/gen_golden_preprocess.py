"""ORACLE -- test infrastructure only.

Generates tests/golden/preprocess.npz by running the REAL reference `seva.eval` preprocessing functions (imported from
/root/reference, build container only) on seeded cases:

    python oracle/gen_golden_preprocess.py

get_resizing_factor (2000 shape pairs, both cover modes), get_wh_with_fixed_shortest_side, transform_img_and_K (crop / pad /
stretch, normalised and pixel intrinsics, size strides, scales) and load_img_and_K on blank frames (torch.Size input).
"""
from __future__ import annotations

import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import refimport  # noqa: E402

TRANSFORM_CASES = [
    # (h, w, size, scale, center, normalised K, size_stride, mode)
    (37, 53, (32, 24), 1.0, (0.5, 0.5), True, 1, "crop"),
    (37, 53, (32, 24), 1.0, (0.5, 0.5), False, 1, "pad"),
    (37, 53, (32, 24), 1.0, (0.5, 0.5), True, 1, "stretch"),
    (64, 48, 40, 1.0, (0.5, 0.5), True, 8, "crop"),
    (48, 64, 40, 1.0, (0.3, 0.6), False, 8, "crop"),
    (90, 120, (72, 72), 1.0, (0.5, 0.5), True, 1, "crop"),
    (90, 120, (72, 72), 1.0, (0.5, 0.5), True, 1, "pad"),
    (120, 90, (64, 96), 1.0, (0.5, 0.5), False, 1, "crop"),
    (120, 90, (64, 96), 1.0, (0.5, 0.5), False, 1, "pad"),
    (100, 100, (50, 50), 0.8, (0.5, 0.5), True, 1, "crop"),
    (75, 131, (64, 32), 1.0, (0.5, 0.5), True, 1, "crop"),
    (144, 256, (128, 128), 1.0, (0.5, 0.5), True, 64, "crop"),
]
LOAD_CASES = [
    # (h, w, size, scale, center, normalised K, size_stride, center_crop)
    (48, 64, (32, 32), 1.0, (0.5, 0.5), True, 1, False),
    (48, 64, 40, 1.0, (0.5, 0.5), False, 8, False),
    (64, 48, (32, 24), 0.8, (0.5, 0.5), True, 1, False),
    (60, 90, (48, 32), 1.0, (0.4, 0.5), True, 1, True),
    (90, 60, None, 1.0, (0.5, 0.5), False, 1, False),
]


def make_K(normalised: bool, w: int, h: int, batched: bool):
    K = torch.tensor([[0.9, 0.0, 0.5], [0.0, 1.1, 0.5], [0.0, 0.0, 1.0]]) if normalised else \
        torch.tensor([[0.9 * w, 0.0, 0.5 * w + 3.0], [0.0, 1.1 * h, 0.5 * h - 2.0], [0.0, 0.0, 1.0]])
    return K[None].clone() if batched else K.clone()


def main():
    ev = refimport.import_eval()
    out = {}
    rng = np.random.RandomState(0)
    shapes = rng.randint(8, 400, size=(2000, 4))
    shapes[:200, 2:] = shapes[:200, :2] * rng.randint(1, 4, size=(200, 1))      # exact multiples: the tie branches
    shapes[200:300, 1] = shapes[200:300, 0]
    shapes[300:400, 3] = shapes[300:400, 2]
    out["rf_shapes"] = shapes
    out["rf_cover"] = np.array([ev.get_resizing_factor((a, b), (c, d), True) for a, b, c, d in shapes.tolist()])
    out["rf_fit"] = np.array([ev.get_resizing_factor((a, b), (c, d), False) for a, b, c, d in shapes.tolist()])
    wh = rng.randint(8, 2000, size=(300, 3))
    out["wh_in"] = wh
    out["wh_out"] = np.array([ev.get_wh_with_fixed_shortest_side(int(a), int(b), int(c)) for a, b, c in wh])
    for i, (h, w, size, scale, center, norm, stride, mode) in enumerate(TRANSFORM_CASES):
        g = torch.Generator().manual_seed(100 + i)
        img = torch.rand(1, 3, h, w, generator=g)
        K = make_K(norm, w, h, True)
        o, K2 = ev.transform_img_and_K(img, size, scale=scale, center=center, K=K, size_stride=stride, mode=mode)
        out[f"t{i}_img"], out[f"t{i}_out"], out[f"t{i}_K"] = img.numpy(), o.numpy(), K2.numpy()
    for i, (h, w, size, scale, center, norm, stride, cc) in enumerate(LOAD_CASES):
        K = make_K(norm, w, h, False)
        o, K2 = ev.load_img_and_K(torch.Size([h, w]), size, scale=scale, center=center, K=K, size_stride=stride,
                                  center_crop=cc, device="cpu")
        out[f"l{i}_out"], out[f"l{i}_K"] = o.numpy(), K2.numpy()
    path = os.path.join(ROOT, "tests", "golden", "preprocess.npz")
    np.savez_compressed(path, **out)
    print("wrote", path, os.path.getsize(path), "bytes")


if __name__ == "__main__":
    main()
