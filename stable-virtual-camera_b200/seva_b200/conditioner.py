"""CLIP image conditioner on the step's kernels behind the reference's ``CLIPConditioner`` surface
(seva/modules/conditioner.py:8-39): ``CLIPConditioner().preprocess(x)`` / ``.forward(x)`` with x ``[N, 3, H, W]`` in [-1, 1]
-> ``[N, 1024]``, the vector ``do_sample`` averages over the input frames and repeats as ``crossattn`` (seva/eval.py:1248).

The reference wraps ``open_clip.create_model_and_transforms("ViT-H-14", pretrained="laion2b_s32b_b79k")[0]`` and calls its
``encode_image``.  open_clip and kornia are absent from this image, so this module holds the visual tower's parameters
itself, under open_clip's own key names (``module.visual.conv1.weight``, ``...transformer.resblocks.{i}.attn.in_proj_weight``,
``...ln_post.weight``, ``module.visual.proj``): ``load_state_dict(open_clip_model_state, strict=False)`` takes a released
checkpoint and ignores the text tower.  Without one the weights are random-initialised (there is no network).

Launch sequence per call (all on the current stream, no PyTorch compute on activations):
  seva_resize_bicubic_aa -> seva_clip_patchify -> patch GEMM (+ positional / class embedding as the residual operand)
  -> ln_pre -> 32 x [LayerNorm, q|k GEMM, per image: V^T GEMM + per head (scores GEMM fp32) + one row softmax + per head
  (P V GEMM), out-proj GEMM + residual, LayerNorm, c_fc GEMM with the erf-GELU epilogue, c_proj GEMM + residual]
  -> ln_post -> class rows -> projection GEMV.
Head width 80 does not fit the flash kernel (64), and at 257 tokens the scores are 0.26 MB per head: attention runs as plain
tensor-core GEMMs around ``seva_softmax_rows`` -- the VAE mid-block's scheme.  Token rows are padded 257 -> 264 per image
(GEMM K operands need multiples of 8); padded keys get probability 0.  GELU reuses the GEGLU epilogue with a constant-1
value half: 1 * gelu(gate).  bf16 activations / weights, fp32 accumulation, statistics and softmax.
No CPU fallback: CPU tensors raise ``SevaB200Error``.
"""
from __future__ import annotations

import math
from dataclasses import dataclass

import torch
from torch import nn

from . import ops, pack
from ._lib import SevaB200Error

BF16, F32 = torch.bfloat16, torch.float32


@dataclass
class ViTParams:
    """open_clip ViT-H-14 visual tower (model_configs/ViT-H-14.json)."""
    image_size: int = 224
    patch_size: int = 14
    width: int = 1280
    layers: int = 32
    heads: int = 16
    mlp_ratio: float = 4.0
    embed_dim: int = 1024


class _Holder(nn.Module):
    """parameter holder: the launch sequence below consumes the parameters; there is no module-level forward"""

    def forward(self, *a, **k):  # pragma: no cover
        raise SevaB200Error("parameter holder: call CLIPConditioner.forward")


def _block(width: int, mlp: int) -> _Holder:
    b = _Holder()
    b.ln_1 = nn.LayerNorm(width)
    b.attn = _Holder()
    b.attn.in_proj_weight = nn.Parameter(torch.empty(3 * width, width))
    b.attn.in_proj_bias = nn.Parameter(torch.zeros(3 * width))
    b.attn.out_proj = nn.Linear(width, width)
    b.ln_2 = nn.LayerNorm(width)
    b.mlp = _Holder()
    b.mlp.c_fc = nn.Linear(width, mlp)
    b.mlp.c_proj = nn.Linear(mlp, width)
    return b


class VisionTower(_Holder):
    def __init__(self, p: ViTParams):
        super().__init__()
        self.p = p
        grid = p.image_size // p.patch_size
        self.conv1 = nn.Conv2d(3, p.width, p.patch_size, p.patch_size, bias=False)
        self.class_embedding = nn.Parameter(torch.empty(p.width))
        self.positional_embedding = nn.Parameter(torch.empty(grid * grid + 1, p.width))
        self.ln_pre = nn.LayerNorm(p.width)
        self.transformer = _Holder()
        self.transformer.resblocks = nn.ModuleList(_block(p.width, int(p.width * p.mlp_ratio)) for _ in range(p.layers))
        self.ln_post = nn.LayerNorm(p.width)
        self.proj = nn.Parameter(torch.empty(p.width, p.embed_dim))


class _ClipModel(_Holder):
    """stands where the reference keeps the open_clip model (``CLIPConditioner.module``)"""

    def __init__(self, p: ViTParams):
        super().__init__()
        self.visual = VisionTower(p)


def _init(t: VisionTower, seed: int) -> None:
    g = torch.Generator().manual_seed(seed)
    w = t.p.width
    with torch.no_grad():
        def rn(x, std):
            x.copy_(torch.randn(x.shape, generator=g) * std)
        rn(t.conv1.weight, (3 * t.p.patch_size ** 2) ** -0.5)
        rn(t.class_embedding, w ** -0.5)
        rn(t.positional_embedding, w ** -0.5)
        rn(t.proj, w ** -0.5)
        for b in t.transformer.resblocks:
            rn(b.attn.in_proj_weight, w ** -0.5)
            rn(b.attn.in_proj_bias, 0.02)
            rn(b.attn.out_proj.weight, w ** -0.5 * (2 * t.p.layers) ** -0.5)
            rn(b.attn.out_proj.bias, 0.02)
            rn(b.mlp.c_fc.weight, w ** -0.5)
            rn(b.mlp.c_fc.bias, 0.02)
            rn(b.mlp.c_proj.weight, (2 * w) ** -0.5 * (2 * t.p.layers) ** -0.5)
            rn(b.mlp.c_proj.bias, 0.02)
        for m in t.modules():
            if isinstance(m, nn.LayerNorm):
                m.weight.copy_(1.0 + 0.1 * torch.randn(m.weight.shape, generator=g))
                m.bias.copy_(0.05 * torch.randn(m.bias.shape, generator=g))


def _f32(t):
    return t.detach().float().contiguous()


def pack_weights(t: VisionTower) -> dict:
    p = t.p
    w, hd = p.width, p.width // p.heads
    if p.width % p.heads or hd % 8 or p.width % 64:
        raise SevaB200Error(f"CLIP tower: width {p.width} / heads {p.heads} unsupported (head width must be a multiple of 8)")
    kin = 3 * p.patch_size ** 2
    kp = (kin + 63) // 64 * 64
    wpatch = torch.zeros((w, kp), dtype=F32, device=t.conv1.weight.device)
    wpatch[:, :kin] = t.conv1.weight.detach().float().reshape(w, kin)
    pos = t.positional_embedding.detach().float().clone()
    pos[0] += t.class_embedding.detach().float()              # row 0: class token (its patch row is zero)
    P = {"wpatch": wpatch.to(BF16).contiguous(), "kp": kp, "pos": pos.to(BF16).contiguous(),
         "ln_pre": (_f32(t.ln_pre.weight), _f32(t.ln_pre.bias)), "ln_post": (_f32(t.ln_post.weight), _f32(t.ln_post.bias)),
         "projT": t.proj.detach().float().t().contiguous().to(BF16), "blocks": []}
    s = 1.0 / math.sqrt(hd)
    mlp = int(p.width * p.mlp_ratio)
    bn = ops.pick_bn(2 * mlp, True)
    for b in t.transformer.resblocks:
        wi, bi = b.attn.in_proj_weight.detach().float(), b.attn.in_proj_bias.detach().float()
        wo, bo = b.attn.out_proj.weight.detach().float(), b.attn.out_proj.bias.detach().float()
        # c_fc -> GELU through the GEGLU epilogue: value half = 0 x + 1, gate half = c_fc
        wg = torch.cat([torch.zeros_like(b.mlp.c_fc.weight.detach().float()), b.mlp.c_fc.weight.detach().float()])
        bg = torch.cat([torch.ones(mlp, device=wg.device), b.mlp.c_fc.bias.detach().float()])
        wg, bg = pack.pack_geglu(wg, bg, bn)
        P["blocks"].append({
            "ln1": (_f32(b.ln_1.weight), _f32(b.ln_1.bias)), "ln2": (_f32(b.ln_2.weight), _f32(b.ln_2.bias)),
            # q pre-scaled by 1 / sqrt(head width); v^T comes out of a GEMM with W_v as the A operand (no per-row bias):
            # rows of softmax sum to one, so b_v moves into the output projection's bias
            "wqk": torch.cat([wi[:w] * s, wi[w:2 * w]]).to(BF16).contiguous(),
            "bqk": torch.cat([bi[:w] * s, bi[w:2 * w]]).contiguous(),
            "wv": wi[2 * w:].to(BF16).contiguous(),
            "wo": wo.to(BF16).contiguous(), "bo": (bo + wo @ bi[2 * w:]).contiguous(),
            "wfc": wg, "bfc": bg, "bn": bn,
            "wproj": pack.pack_linear(b.mlp.c_proj.weight), "bproj": _f32(b.mlp.c_proj.bias)})
    return P


def _encode_cuda(P: dict, p: ViTParams, mean, std, x: torch.Tensor) -> torch.Tensor:
    dev = x.device
    N, w, H, hd = x.shape[0], p.width, p.heads, p.width // p.heads
    G = p.image_size // p.patch_size
    L = G * G + 1
    Lp = (L + 7) // 8 * 8
    A = ops.clip_patchify(x, p.patch_size, Lp, P["kp"], mean, std)              # [N Lp, kp]
    posadd = torch.zeros((N, Lp, w), device=dev, dtype=BF16)
    posadd[:, :L] = P["pos"]                                                     # plumbing copy of a parameter
    posadd = posadd.view(N * Lp, w)
    tok = ops.linear(A, P["wpatch"], None, residual=posadd)
    xs = ops.layernorm(tok, *P["ln_pre"], eps=1e-5)
    S = torch.empty((H * Lp, Lp), device=dev, dtype=F32)
    Pr = torch.zeros((H * Lp, Lp), device=dev, dtype=BF16)                       # padded key columns stay 0
    vt = torch.empty((w, Lp), device=dev, dtype=BF16)
    o = torch.empty((N * Lp, w), device=dev, dtype=BF16)
    for blk in P["blocks"]:
        h = ops.layernorm(xs, *blk["ln1"], eps=1e-5)
        qk = ops.linear(h, blk["wqk"], blk["bqk"])                               # [N Lp, 2 w]  (q pre-scaled | k)
        for i in range(N):
            r = slice(i * Lp, (i + 1) * Lp)
            ops.linear(blk["wv"], h[r], None, out=vt)                            # v^T = W_v h^T  [w, Lp]
            for hh in range(H):
                c = slice(hh * hd, (hh + 1) * hd)
                ops.linear(qk[r, c], qk[r, w + hh * hd:w + (hh + 1) * hd], None, out=S[hh * Lp:(hh + 1) * Lp],
                           out_fp32=True)
            ops.softmax_rows(S[:, :L], out=Pr[:, :L])
            for hh in range(H):
                ops.linear(Pr[hh * Lp:(hh + 1) * Lp], vt[hh * hd:(hh + 1) * hd], None, out=o[r, hh * hd:(hh + 1) * hd])
        xs = ops.linear(o, blk["wo"], blk["bo"], residual=xs)
        h = ops.layernorm(xs, *blk["ln2"], eps=1e-5)
        m = ops.linear(h, blk["wfc"], blk["bfc"], geglu=True, bn=blk["bn"])
        xs = ops.linear(m, blk["wproj"], blk["bproj"], residual=xs)
    xs = ops.layernorm(xs, *P["ln_post"], eps=1e-5)
    cls = xs.view(N, Lp, w)[:, 0].float().contiguous()                           # [N, w] class rows (plumbing copy)
    return ops.gemv_rows(cls, P["projT"], None)


class CLIPConditioner(nn.Module):
    """Drop-in for seva.modules.conditioner.CLIPConditioner (conditioner.py:8-39)."""
    mean: torch.Tensor
    std: torch.Tensor

    def __init__(self, params: ViTParams | None = None, state_dict: dict | None = None, seed: int | None = 0):
        super().__init__()
        self.params = params or ViTParams()
        self.module = _ClipModel(self.params)
        if state_dict is not None:
            self.module.load_state_dict(state_dict, strict=False)
        elif seed is not None:
            _init(self.module.visual, seed)
        self.module.eval().requires_grad_(False)
        self.register_buffer("mean", torch.Tensor([0.48145466, 0.4578275, 0.40821073]), persistent=False)
        self.register_buffer("std", torch.Tensor([0.26862954, 0.26130258, 0.27577711]), persistent=False)
        self._packed = None

    def invalidate(self) -> None:
        self._packed = None

    def _apply(self, fn, *a, **k):
        self._packed = None
        return super()._apply(fn, *a, **k)

    def load_state_dict(self, sd, strict: bool = True, **kw):
        self._packed = None
        return super().load_state_dict(sd, strict=strict, **kw)

    def _weights(self) -> dict:
        dev = self.module.visual.conv1.weight.device
        if dev.type != "cuda":
            raise SevaB200Error("CLIPConditioner: parameters are not on a CUDA device (seva_b200 has no CPU fallback)")
        if self._packed is None:
            with torch.cuda.device(dev):
                self._packed = pack_weights(self.module.visual)
        return self._packed

    def _check(self, x: torch.Tensor) -> None:
        if not x.is_cuda:
            raise SevaB200Error("CLIPConditioner: input must be a CUDA tensor (seva_b200 has no CPU fallback)")
        if x.dim() != 4 or x.shape[1] != 3:
            raise ValueError(f"CLIPConditioner: expected [N, 3, H, W], got {tuple(x.shape)}")

    def preprocess(self, x: torch.Tensor) -> torch.Tensor:
        """conditioner.py:24-35: bicubic antialiased resize to 224 x 224, (x + 1) / 2, mean / std normalisation."""
        self._check(x)
        s = self.params.image_size
        with torch.cuda.device(x.device):
            y = ops.resize_bicubic_aa(x.float(), (s, s), antialias=True)
            mean = self.mean.to(y).view(1, 3, 1, 1)
            std = self.std.to(y).view(1, 3, 1, 1)
            return ((y + 1.0) / 2.0 - mean) / std        # API-parity helper only; forward() normalises inside seva_clip_patchify

    @torch.no_grad()
    def forward(self, x: torch.Tensor) -> torch.Tensor:
        """conditioner.py:37-39: preprocess + open_clip ``encode_image`` -> [N, embed_dim] fp32."""
        self._check(x)
        P = self._weights()
        s = self.params.image_size
        with torch.cuda.device(x.device):
            y = ops.resize_bicubic_aa(x.float(), (s, s), antialias=True)
            return _encode_cuda(P, self.params, self.mean.tolist(), self.std.tolist(), y)
