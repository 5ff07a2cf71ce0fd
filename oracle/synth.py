"""ORACLE — test infrastructure only.

Deterministic synthetic weights and inputs shared by the golden-vector generator (which
feeds them to the real reference) and by the tests (which feed them to the oracle restatement and to the CUDA
path).  Everything is drawn from seeded CPU generators, so the same tensors are rebuilt on any machine.
"""
from __future__ import annotations

import math
import zlib

import torch


def fill_state_dict(shapes: dict, seed: int = 0) -> dict:
    """name -> fp32 tensor.  Matrices/filters ~ U(-1,1)/sqrt(fan_in) (PyTorch's default scale), norm gains
    1 + 0.1 N(0,1), biases 0.05 N(0,1); each tensor has its own generator seeded from (seed, crc32(name))."""
    out = {}
    for name in sorted(shapes):
        shape = tuple(shapes[name])
        g = torch.Generator().manual_seed((seed * 1000003 + zlib.crc32(name.encode())) % (2**63 - 1))
        if len(shape) >= 2:
            fan_in = math.prod(shape[1:])
            t = (torch.rand(shape, generator=g) * 2 - 1) / math.sqrt(fan_in)
        elif name.endswith("weight"):
            t = 1.0 + 0.1 * torch.randn(shape, generator=g)
        else:
            t = 0.05 * torch.randn(shape, generator=g)
        out[name] = t
    return out


def orbit_cameras(T: int, radius: float = 2.0, arc: float = 0.25) -> torch.Tensor:
    """c2w [T,4,4] (OpenCV convention: +z forward) on a circle of ``radius`` looking at the origin,
    yaw_i = 2 pi * arc * i / T."""
    c2ws = []
    for i in range(T):
        yaw = 2 * math.pi * arc * i / T
        pos = torch.tensor([radius * math.sin(yaw), 0.0, -radius * math.cos(yaw)])
        fwd = -pos / pos.norm()
        up = torch.tensor([0.0, -1.0, 0.0])
        right = torch.linalg.cross(fwd, up)
        right = right / right.norm()
        down = torch.linalg.cross(fwd, right)
        m = torch.eye(4)
        m[:3, 0], m[:3, 1], m[:3, 2], m[:3, 3] = right, down, fwd, pos
        c2ws.append(m)
    return torch.stack(c2ws)


def make_inputs(T: int, n_input: int, h: int, w: int, plucker_fn, seed: int = 1, num_steps: int = 1,
                ctx_dim: int = 1024) -> dict:
    """Sampling-run inputs built exactly as eval.py:1249-1281 does from (latents, CLIP token, Plucker map).

    plucker_fn(w2c_src, w2c, K_norm, h, w) -> [T,6,h,w] supplies the ray map (reference, oracle or CUDA).
    """
    g = torch.Generator().manual_seed(seed)
    c2w = orbit_cameras(T)
    w2c = torch.linalg.inv(c2w)
    f = 0.5 / math.tan(0.5 * 0.9424777960769379)
    K = torch.tensor([[f, 0.0, 0.5], [0.0, f, 0.5], [0.0, 0.0, 1.0]]).repeat(T, 1, 1)
    mask = torch.zeros(T, dtype=torch.bool)
    mask[:n_input] = True
    pl = plucker_fn(w2c[0], w2c, K.clone(), h, w)
    latents = torch.randn(n_input, 4, h, w, generator=g)
    clip = torch.randn(ctx_dim, generator=g)
    x0 = torch.randn(T, 4, h, w, generator=g)
    eps = torch.randn(num_steps, T, 4, h, w, generator=g)
    c_cross = clip[None, None].repeat(T, 1, 1)
    c_replace = torch.zeros(T, 5, h, w)
    c_replace[mask] = torch.nn.functional.pad(latents, (0, 0, 0, 0, 0, 1), value=1.0)
    c_concat = torch.cat([mask[:, None, None, None].float().expand(T, 1, h, w), pl], 1)
    uc_concat = torch.cat([torch.zeros(T, 1, h, w), pl], 1)
    c = {"crossattn": c_cross, "replace": c_replace, "concat": c_concat, "dense_vector": pl}
    uc = {"crossattn": torch.zeros_like(c_cross), "replace": torch.zeros_like(c_replace), "concat": uc_concat,
          "dense_vector": pl}
    return dict(c2w=c2w, w2c=w2c, K=K, mask=mask, plucker=pl, c=c, uc=uc, x0=x0, eps=eps)
