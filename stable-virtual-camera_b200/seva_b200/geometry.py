"""Camera geometry on the hot path, behind the reference's ``seva.geometry`` names.

Mirrors seva/geometry.py:9-165: ``DEFAULT_FOV_RAD``, ``get_camera_dist`` (:12-40), ``to_hom_pose`` (:49-55),
``get_default_intrinsics`` (:58-79) and ``get_plucker_coordinates`` (:119-165).  The ray map itself is one CUDA
kernel (``seva_plucker``); the intrinsics bookkeeping (normalisation check and the in-place scaling of the
caller's tensor, geometry.py:133,149) is kept as the reference does it because callers rely on it.
"""
from __future__ import annotations

import torch

from . import ops
from ._lib import SevaB200Error

DEFAULT_FOV_RAD = 0.9424777960769379  # 54 degrees


def get_camera_dist(source_c2ws: torch.Tensor, target_c2ws: torch.Tensor, mode: str = "translation") -> torch.Tensor:
    """Pairwise camera distance [N, M]: rotation angle in degrees or translation L2 (geometry.py:12-40)."""
    if mode == "rotation":
        ra, rb = source_c2ws[:, None, :3, :3], target_c2ws[None, :, :3, :3]
        trace = (ra * rb).sum(dim=(-1, -2))  # tr(Ra Rb^T)
        return torch.acos(((trace - 1) / 2).clamp(-1, 1)) * (180 / torch.pi)
    if mode == "translation":
        return torch.norm(source_c2ws[:, None, :3, 3] - target_c2ws[None, :, :3, 3], dim=-1)
    raise NotImplementedError(f"Mode {mode} is not implemented for finding nearest source indices.")


def to_hom_pose(pose: torch.Tensor) -> torch.Tensor:
    if pose.shape[-2:] == (3, 4):
        hom = torch.eye(4, device=pose.device)[None].repeat(pose.shape[0], 1, 1)
        hom[:, :3, :] = pose
        return hom
    return pose


def get_default_intrinsics(fov_rad=DEFAULT_FOV_RAD, aspect_ratio=1.0) -> torch.Tensor:
    if not isinstance(fov_rad, torch.Tensor):
        fov_rad = torch.tensor([fov_rad] if isinstance(fov_rad, (int, float)) else fov_rad)
    if aspect_ratio >= 1.0:  # W >= H
        focal_x = 0.5 / torch.tan(0.5 * fov_rad)
        focal_y = focal_x * aspect_ratio
    else:
        focal_y = 0.5 / torch.tan(0.5 * fov_rad)
        focal_x = focal_y / aspect_ratio
    K = focal_x.new_zeros((focal_x.shape[0], 3, 3))
    K[:, 0, 0], K[:, 1, 1], K[:, 2, 2] = focal_x, focal_y, 1.0
    K[:, 0, 2], K[:, 1, 2] = 0.5, 0.5
    return K


def get_plucker_coordinates(extrinsics_src, extrinsics, intrinsics=None, fov_rad=DEFAULT_FOV_RAD,
                            target_size=[72, 72]) -> torch.Tensor:
    """[T, 6, h, w] Plucker rays of every camera relative to ``extrinsics_src`` (both world-to-camera 4x4)."""
    if intrinsics is None:
        intrinsics = get_default_intrinsics(fov_rad).to(extrinsics.device)
    else:
        pp = intrinsics[:, :2, -1]
        if not (torch.all(pp >= 0) and torch.all(pp <= 1)):
            intrinsics[:, :2] /= intrinsics.new_tensor(target_size).view(1, -1, 1) * 8
        pp = intrinsics[:, :2, -1]
        assert torch.all(pp >= 0) and torch.all(pp <= 1), \
            "Intrinsics should be expressed in resolution-independent normalized image coordinates."
    # in place, like the reference: callers pass a clone (eval.py:1206)
    intrinsics[:, :2] *= extrinsics.new_tensor([target_size[1], target_size[0]]).view(1, -1, 1)

    if not torch.cuda.is_available():
        raise SevaB200Error("get_plucker_coordinates needs a CUDA device (seva_b200 has no CPU fallback)")
    src_dev = extrinsics.device
    dev = src_dev if src_dev.type == "cuda" else torch.device("cuda", torch.cuda.current_device())
    T = extrinsics.shape[0]
    w2c = to_hom_pose(extrinsics.float()).to(dev).contiguous()
    w2c_src = extrinsics_src.float().to(dev)
    if w2c_src.shape[-2:] == (3, 4):
        w2c_src = to_hom_pose(w2c_src[None])[0]
    Kpix = intrinsics.float().to(dev)
    if Kpix.shape[0] != T:
        Kpix = Kpix.expand(T, 3, 3)
    with torch.cuda.device(dev):
        out = ops.plucker(w2c_src.contiguous(), w2c, Kpix.contiguous(), int(target_size[0]), int(target_size[1]))
    return out.to(src_dev)
