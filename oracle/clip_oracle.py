"""TEST INFRASTRUCTURE ONLY -- fp32 PyTorch restatement of the CLIP conditioner.

Only tests/ may import this module; the product path (seva_b200/conditioner.py) never does.

PARITY UNPINNED: the reference's CLIPConditioner (seva/modules/conditioner.py:1-39) is a thin wrapper over two third-party
packages that are absent from this image -- open_clip (ViT-H-14, `laion2b_s32b_b79k`; unpinned in the reference's
pyproject.toml) and kornia (unpinned) -- so no reference output can be generated here.  The functions below restate their
published algorithms:
  * kornia.geometry.transform.resize(..., interpolation="bicubic", align_corners=True, antialias=True): when the larger of
    the two downscale factors exceeds 1, gaussian_blur2d(kernel_size = int(max(4 sigma, 3)) made odd, sigma =
    max((factor - 1) / 2, 0.001), border_type="reflect", separable) and then torch.nn.functional.interpolate(bicubic);
  * kornia.enhance.normalize: (x - mean) / std per channel;
  * open_clip.transformer.VisionTransformer.forward (pool_type "tok", no attentional pooling): conv1 (P x P, stride P, no
    bias) -> [class_embedding; patches] + positional_embedding -> ln_pre -> `layers` ResidualAttentionBlocks
    (x += attn(ln_1 x); x += c_proj(gelu(c_fc(ln_2 x))), nn.MultiheadAttention with in_proj / out_proj, exact erf GELU)
    -> ln_post -> class token -> @ proj.
`vit_forward` is checked against an independent implementation of the same architecture, Hugging Face transformers'
CLIPVisionModelWithProjection with the weights converted key by key (tests/test_clip_hf_cpu.py, rel-L2 < 2e-5); the
resize stays a restatement.  The torch pieces the restatement leans on (F.interpolate, F.layer_norm, F.gelu, softmax) are the library calls the real
packages make themselves.
"""
from __future__ import annotations

import math

import torch
import torch.nn.functional as F

CLIP_MEAN = (0.48145466, 0.4578275, 0.40821073)   # seva/modules/conditioner.py:17-19
CLIP_STD = (0.26862954, 0.26130258, 0.27577711)   # :20-22


def gaussian_kernel1d(ks: int, sigma: float) -> torch.Tensor:
    x = torch.arange(ks, dtype=torch.float64) - ks // 2
    g = torch.exp(-(x * x) / (2.0 * sigma * sigma))
    return (g / g.sum()).float()


def blur_params(in_size: int, out_size: int):
    factor = in_size / out_size
    sigma = max((factor - 1.0) / 2.0, 0.001)
    ks = int(max(2.0 * 2 * sigma, 3))
    if ks % 2 == 0:
        ks += 1
    return ks, sigma


def resize_bicubic_aa(x: torch.Tensor, size: tuple[int, int], antialias: bool = True) -> torch.Tensor:
    """kornia.geometry.resize(x, size, interpolation="bicubic", align_corners=True, antialias=antialias); x [N, C, H, W]."""
    H, W = x.shape[-2:]
    fy, fx = H / size[0], W / size[1]
    if antialias and max(fy, fx) > 1:
        (ky, sy), (kx, sx) = blur_params(H, size[0]), blur_params(W, size[1])
        C = x.shape[1]
        gy = gaussian_kernel1d(ky, sy).to(x)
        gx = gaussian_kernel1d(kx, sx).to(x)
        xp = F.pad(x, (kx // 2, kx // 2, ky // 2, ky // 2), mode="reflect")
        xp = F.conv2d(xp, gx.view(1, 1, 1, kx).expand(C, 1, 1, kx), groups=C)
        x = F.conv2d(xp, gy.view(1, 1, ky, 1).expand(C, 1, ky, 1), groups=C)
    return F.interpolate(x, size=size, mode="bicubic", align_corners=True)


def preprocess(x: torch.Tensor, size: int = 224) -> torch.Tensor:
    """CLIPConditioner.preprocess (conditioner.py:24-35): x in [-1, 1] -> normalised [N, 3, size, size]."""
    x = resize_bicubic_aa(x, (size, size))
    x = (x + 1.0) / 2.0
    mean = torch.tensor(CLIP_MEAN, dtype=x.dtype, device=x.device).view(1, 3, 1, 1)
    std = torch.tensor(CLIP_STD, dtype=x.dtype, device=x.device).view(1, 3, 1, 1)
    return (x - mean) / std


def _strip(sd: dict) -> dict:
    """accept `visual.*`, `module.visual.*` or bare keys"""
    out = {}
    for k, v in sd.items():
        for pre in ("module.visual.", "visual."):
            if k.startswith(pre):
                k = k[len(pre):]
                break
        out[k] = v.float()
    return out


def vit_forward(sd: dict, x: torch.Tensor, heads: int) -> torch.Tensor:
    """open_clip VisionTransformer.forward on a preprocessed batch x [N, 3, S, S] -> [N, embed_dim]."""
    p = _strip(sd)
    w = p["conv1.weight"]
    width, patch = w.shape[0], w.shape[-1]
    x = F.conv2d(x, w, stride=patch)                                   # [N, width, G, G]
    N = x.shape[0]
    x = x.reshape(N, width, -1).permute(0, 2, 1)                       # [N, G*G, width]
    x = torch.cat([p["class_embedding"].view(1, 1, -1).expand(N, -1, -1), x], dim=1)
    x = x + p["positional_embedding"]
    x = F.layer_norm(x, (width,), p["ln_pre.weight"], p["ln_pre.bias"], 1e-5)
    hd = width // heads
    i = 0
    while f"transformer.resblocks.{i}.ln_1.weight" in p:
        b = f"transformer.resblocks.{i}."
        h = F.layer_norm(x, (width,), p[b + "ln_1.weight"], p[b + "ln_1.bias"], 1e-5)
        qkv = F.linear(h, p[b + "attn.in_proj_weight"], p[b + "attn.in_proj_bias"])
        q, k, v = (t.reshape(N, -1, heads, hd).transpose(1, 2) for t in qkv.chunk(3, dim=-1))
        a = torch.softmax((q @ k.transpose(-1, -2)) / math.sqrt(hd), dim=-1) @ v
        a = a.transpose(1, 2).reshape(N, -1, width)
        x = x + F.linear(a, p[b + "attn.out_proj.weight"], p[b + "attn.out_proj.bias"])
        h = F.layer_norm(x, (width,), p[b + "ln_2.weight"], p[b + "ln_2.bias"], 1e-5)
        h = F.gelu(F.linear(h, p[b + "mlp.c_fc.weight"], p[b + "mlp.c_fc.bias"]))
        x = x + F.linear(h, p[b + "mlp.c_proj.weight"], p[b + "mlp.c_proj.bias"])
        i += 1
    x = F.layer_norm(x, (width,), p["ln_post.weight"], p["ln_post.bias"], 1e-5)
    return x[:, 0] @ p["proj"]


def conditioner_forward(sd: dict, x: torch.Tensor, heads: int, size: int = 224) -> torch.Tensor:
    """CLIPConditioner.forward (conditioner.py:37-39): preprocess + encode_image."""
    return vit_forward(sd, preprocess(x.float(), size), heads)


# ---- seva/eval.py:249-322 transform_img_and_K: only F.interpolate(mode="area") is arithmetic; restated for the kernel test ----
def resize_area(x: torch.Tensor, size: tuple[int, int]) -> torch.Tensor:
    return F.interpolate(x, size, mode="area", antialias=False)
