"""Input-frame preprocessing of the scene loop: the resize / crop / pad of ``load_img_and_K`` and
``transform_img_and_K`` (seva/eval.py:160-322) together with their intrinsics bookkeeping.  Same names, argument meaning,
return values and assertion behaviour as the reference; the one arithmetic step -- ``F.interpolate(mode="area")`` -- runs as
``seva_resize_area`` on the GPU, everything else is integer / 3x3 bookkeeping on the host plus slicing.

Images are CUDA tensors here (the reference resizes on the CPU and moves the result, eval.py:241): a CPU image raises
``SevaB200Error`` -- there is no CPU fallback.  ``K`` stays wherever the caller keeps it.
"""
from __future__ import annotations

import math
from typing import Optional, Tuple, Union

import numpy as np
import torch
import torch.nn.functional as F

from . import ops
from ._lib import SevaB200Error


def get_resizing_factor(target_shape: Tuple[int, int], current_shape: Tuple[int, int], cover_target: bool = True) -> float:
    """eval.py:99-138.  Scale factor that makes `current_shape` (H, W) cover the target (cover_target) or fit inside it.

    The reference walks a 12-way table over the two aspect ratios; every leaf is one of the two quotients below, chosen by
    whether the current shape is relatively wider than the target (ties go the way the table sends them)."""
    th, tw = target_shape
    ch, cw = current_shape
    r_bound, aspect = tw / th, cw / ch
    by_h, by_w = th / ch, tw / cw
    if r_bound >= 1.0:
        wider = aspect >= r_bound
        return (by_h if wider else by_w) if cover_target else (by_w if wider else by_h)
    narrower = aspect <= r_bound
    return (by_w if narrower else by_h) if cover_target else (by_h if narrower else by_w)


def get_wh_with_fixed_shortest_side(w: int, h: int, size: Optional[int]):
    """eval.py:147-157."""
    if size is None or size <= 0:
        return w, h
    if w < h:
        return size, int(size * h / w)
    return int(size * w / h), size


def _target_wh(w: int, h: int, size, size_stride: int):
    if isinstance(size, (tuple, list)):
        W, H = size
    else:
        W, H = get_wh_with_fixed_shortest_side(w, h, size)
    return (math.floor(W / size_stride + 0.5) * size_stride, math.floor(H / size_stride + 0.5) * size_stride)


def _area(image: torch.Tensor, rh: int, rw: int) -> torch.Tensor:
    if not image.is_cuda:
        raise SevaB200Error("preprocess: the image must be a CUDA tensor (seva_b200 has no CPU fallback)")
    return ops.resize_area(image.float(), (rh, rw))


def _scale_K(K: torch.Tensor, rw: int, rh: int, w: int, h: int, batched: bool) -> torch.Tensor:
    K = K.clone()
    pp = K[:, :2, -1] if batched else K[:2, -1]
    rows = K[:, :2] if batched else K[:2]
    if torch.all(pp >= 0) and torch.all(pp <= 1):
        f = K.new_tensor([rw, rh])            # normalized K
    else:
        f = K.new_tensor([rw / w, rh / h])    # unnormalized K
    rows *= f[None, :, None] if batched else f[:, None]
    return K


def transform_img_and_K(image: torch.Tensor, size: Union[int, Tuple[int, int]], scale: float = 1.0,
                        center: Tuple[float, float] = (0.5, 0.5), K: torch.Tensor | None = None, size_stride: int = 1,
                        mode: str = "crop"):
    """eval.py:249-319: area-resize `image` [..., h, w] and crop / pad it to `size`; K [B, 3, 3] follows."""
    assert mode in ["crop", "pad", "stretch"], f"mode should be one of ['crop', 'pad', 'stretch'], got {mode}"
    h, w = image.shape[-2:]
    W, H = _target_wh(w, h, size, size_stride)
    if mode == "stretch":
        rh, rw = H, W
    else:
        rfs = get_resizing_factor((H, W), (h, w), cover_target=mode != "pad")
        rh, rw = (int(np.ceil(rfs * s)) for s in (h, w))
    rh, rw = int(rh / scale), int(rw / scale)
    image = _area(image, rh, rw)
    cy_center, cx_center = int(center[1] * rh), int(center[0] * rw)
    if mode != "pad":
        ct = min(max(0, cy_center - H // 2), rh - H)
        cl = min(max(0, cx_center - W // 2), rw - W)
        image = image[..., ct:ct + H, cl:cl + W].contiguous()
        pl, pt = 0, 0
    else:
        pt, pl = max(0, H // 2 - cy_center), max(0, W // 2 - cx_center)
        pb, pr = max(0, H - pt - rh), max(0, W - pl - rw)
        image = F.pad(image, [pl, pr, pt, pb])
        cl, ct = 0, 0
    if K is not None:
        K = _scale_K(K, rw, rh, w, h, batched=True)
        K[:, :2, 2] += K.new_tensor([pl - cl, pt - ct])
    return image, K


def load_img_and_K(image_path_or_size: Union[str, torch.Size], size: Optional[Union[int, Tuple[int, int]]],
                   scale: float = 1.0, center: Tuple[float, float] = (0.5, 0.5), K: torch.Tensor | None = None,
                   size_stride: int = 1, center_crop: bool = False, image_as_tensor: bool = True,
                   context_rgb: np.ndarray | None = None, device: str = "cuda"):
    """eval.py:160-246: decode (or blank) an RGBA frame, composite on white / `context_rgb`, area-resize, pad, crop.

    Returns the (1, 3, H, W) image in [-1, 1] on `device` (or a PIL image when ``image_as_tensor`` is false) and K [3, 3]."""
    from PIL import Image

    if isinstance(image_path_or_size, torch.Size):
        image = Image.new("RGBA", image_path_or_size[::-1])
    else:
        image = Image.open(image_path_or_size).convert("RGBA")
    w, h = image.size
    if size is None:
        size = (w, h)
    arr = np.array(image).astype(np.float32) / 255
    if arr.shape[-1] == 4:
        rgb, alpha = arr[:, :, :3], arr[:, :, 3:]
        arr = rgb * alpha + (context_rgb if context_rgb is not None else 1.0) * (1 - alpha)
    dev = torch.device(device)
    if dev.type != "cuda":
        raise SevaB200Error("load_img_and_K: device must be a CUDA device (seva_b200 has no CPU fallback)")
    img = torch.from_numpy(np.ascontiguousarray(arr.transpose(2, 0, 1))).to(dtype=torch.float32).unsqueeze(0).to(dev)
    W, H = _target_wh(w, h, size, size_stride)
    rfs = get_resizing_factor((math.floor(H * scale), math.floor(W * scale)), (h, w))
    rh, rw = (int(np.ceil(rfs * s)) for s in (h, w))
    img = _area(img, rh, rw)
    if scale < 1.0:
        pw, ph = math.ceil((W - rw) * 0.5), math.ceil((H - rh) * 0.5)
        img = F.pad(img, (pw, pw, ph, ph), "constant", 1.0)
    ih, iw = img.shape[-2:]
    cy_center, cx_center = int(center[1] * ih), int(center[0] * iw)
    ch, cw = (min(H, W),) * 2 if center_crop else (H, W)
    ct = min(max(0, cy_center - ch // 2), ih - ch)
    cl = min(max(0, cx_center - cw // 2), iw - cw)
    img = img[..., ct:ct + ch, cl:cl + cw].contiguous()
    if K is not None:
        K = _scale_K(K, rw, rh, w, h, batched=False)
        K[:2, 2] -= K.new_tensor([cl, ct])
    if image_as_tensor:
        img = img * 2.0 - 1.0
    else:
        img = Image.fromarray((img.permute(0, 2, 3, 1).cpu().numpy()[0] * 255).astype(np.uint8))
    return img, K
