"""The VAE either side of the sampling loop on B200 kernels, behind the reference's ``seva.modules.autoencoder`` surface.

Mirrors ``AutoEncoder`` (seva/modules/autoencoder.py:6-51): ``scale_factor`` / ``downsample``, ``encode`` / ``decode``
with the frame ``chunk_size`` of the reference (:27-36, :41-50), ``_encode`` = ``latent_dist.mean * scale_factor``
(:21-25), ``_decode`` = ``decode(z / scale_factor).sample`` (:38-39), ``forward`` = decode(encode(x)) (:52).  The wrapped
network is diffusers' ``AutoencoderKL`` in the ``stabilityai/stable-diffusion-2-1-base`` configuration (3 -> 128 / 256 /
512 / 512 -> 4 latent channels, 2 ResNet layers per block, GroupNorm32 eps 1e-6, one-head mid-block attention);
``self.module`` holds its parameters under diffusers' names, so a diffusers state dict loads unchanged (the pre-0.19
attention names ``query / key / value / proj_attn`` are accepted too).  diffusers itself is absent from this image and
the pretrained weights need the network: the constructor therefore does NOT download -- it starts from random-init
parameters (``seed``) or from a state dict handed in by the caller.

Execution: the kernels of the denoising step -- 3x3 convolutions as implicit GEMMs on
tcgen05 over the zero-bordered channels-last layout, GroupNorm(+SiLU), nearest 2x upsampling, the stride-2 gather -- plus
a row softmax for the single 512-wide attention head, whose scores and P.V products are plain GEMMs.  bf16 activations,
fp32 accumulation / statistics / softmax.  No CPU fallback: CPU tensors raise ``SevaB200Error``.
"""
from __future__ import annotations

import math

import torch
import torch.nn as nn

from . import ops, pack
from ._lib import SevaB200Error

BF16, F32 = torch.bfloat16, torch.float32

BLOCK_OUT = (128, 256, 512, 512)
LAYERS_PER_BLOCK = 2
LATENT = 4
GROUPS = 32
EPS = 1e-6
_CPAD = 64   # narrow inputs (RGB, latents) are padded to one 64-channel K slab, like the step's conv_in

_OLD_ATTN = {"query": "to_q", "key": "to_k", "value": "to_v", "proj_attn": "to_out.0"}


class _Holder(nn.Module):
    """Parameter container (nested by name): compute happens in the launch sequence below."""

    def forward(self, *a, **k):  # pragma: no cover
        raise SevaB200Error("AutoencoderKL holds parameters only; call AutoEncoder.encode / decode")


def _resnet(cin: int, cout: int) -> _Holder:
    m = _Holder()
    m.norm1 = nn.GroupNorm(GROUPS, cin, eps=EPS)
    m.conv1 = nn.Conv2d(cin, cout, 3, 1, 1)
    m.norm2 = nn.GroupNorm(GROUPS, cout, eps=EPS)
    m.conv2 = nn.Conv2d(cout, cout, 3, 1, 1)
    if cin != cout:
        m.conv_shortcut = nn.Conv2d(cin, cout, 1)
    return m


def _mid(c: int) -> _Holder:
    m = _Holder()
    att = _Holder()
    att.group_norm = nn.GroupNorm(GROUPS, c, eps=EPS)
    att.to_q, att.to_k, att.to_v = nn.Linear(c, c), nn.Linear(c, c), nn.Linear(c, c)
    att.to_out = nn.ModuleList([nn.Linear(c, c)])
    m.attentions = nn.ModuleList([att])
    m.resnets = nn.ModuleList([_resnet(c, c), _resnet(c, c)])
    return m


def _sampler(c: int) -> _Holder:
    m = _Holder()
    m.conv = nn.Conv2d(c, c, 3)
    return m


class AutoencoderKL(_Holder):
    """Parameter tree with diffusers' ``AutoencoderKL`` names (SD-2.1 VAE configuration)."""

    def __init__(self):
        super().__init__()
        enc = _Holder()
        enc.conv_in = nn.Conv2d(3, BLOCK_OUT[0], 3, 1, 1)
        enc.down_blocks = nn.ModuleList()
        cin = BLOCK_OUT[0]
        for i, cout in enumerate(BLOCK_OUT):
            blk = _Holder()
            blk.resnets = nn.ModuleList([_resnet(cin if j == 0 else cout, cout) for j in range(LAYERS_PER_BLOCK)])
            if i != len(BLOCK_OUT) - 1:
                blk.downsamplers = nn.ModuleList([_sampler(cout)])
            enc.down_blocks.append(blk)
            cin = cout
        enc.mid_block = _mid(BLOCK_OUT[-1])
        enc.conv_norm_out = nn.GroupNorm(GROUPS, BLOCK_OUT[-1], eps=EPS)
        enc.conv_out = nn.Conv2d(BLOCK_OUT[-1], 2 * LATENT, 3, 1, 1)
        self.encoder = enc

        dec = _Holder()
        rev = tuple(reversed(BLOCK_OUT))
        dec.conv_in = nn.Conv2d(LATENT, rev[0], 3, 1, 1)
        dec.mid_block = _mid(rev[0])
        dec.up_blocks = nn.ModuleList()
        cin = rev[0]
        for i, cout in enumerate(rev):
            blk = _Holder()
            blk.resnets = nn.ModuleList([_resnet(cin if j == 0 else cout, cout) for j in range(LAYERS_PER_BLOCK + 1)])
            if i != len(rev) - 1:
                blk.upsamplers = nn.ModuleList([_sampler(cout)])
            dec.up_blocks.append(blk)
            cin = cout
        dec.conv_norm_out = nn.GroupNorm(GROUPS, rev[-1], eps=EPS)
        dec.conv_out = nn.Conv2d(rev[-1], 3, 3, 1, 1)
        self.decoder = dec
        self.quant_conv = nn.Conv2d(2 * LATENT, 2 * LATENT, 1)
        self.post_quant_conv = nn.Conv2d(LATENT, LATENT, 1)


# ---- weight packing ------------------------------------------------------------------------------------------------
def _f32(t):
    return t.detach().float().contiguous()


def _pack_resnet(m) -> dict:
    d = {"g1": _f32(m.norm1.weight), "b1": _f32(m.norm1.bias), "w1": pack.pack_conv3x3(m.conv1.weight),
         "c1": _f32(m.conv1.bias), "g2": _f32(m.norm2.weight), "b2": _f32(m.norm2.bias),
         "w2": pack.pack_conv3x3(m.conv2.weight), "c2": _f32(m.conv2.bias), "ws": None, "cs": None}
    if hasattr(m, "conv_shortcut"):
        d["ws"], d["cs"] = pack.pack_conv1x1(m.conv_shortcut.weight), _f32(m.conv_shortcut.bias)
    return d


def _pack_mid(m) -> dict:
    a = m.attentions[0]
    c = a.to_q.weight.shape[0]
    s = 1.0 / math.sqrt(c)       # one head of width c (diffusers Attention: heads = 1, dim_head = 512); folded into q
    wv, bv = a.to_v.weight.detach().float(), a.to_v.bias.detach().float()
    wo, bo = a.to_out[0].weight.detach().float(), a.to_out[0].bias.detach().float()
    return {"r0": _pack_resnet(m.resnets[0]), "r1": _pack_resnet(m.resnets[1]),
            "g": _f32(a.group_norm.weight), "b": _f32(a.group_norm.bias),
            "wqk": torch.cat([a.to_q.weight.detach().float() * s, a.to_k.weight.detach().float()]).to(BF16).contiguous(),
            "bqk": torch.cat([a.to_q.bias.detach().float() * s, a.to_k.bias.detach().float()]).contiguous(),
            # v^T comes out of a GEMM with the weight as the A operand, which has no per-row bias: rows of softmax sum to
            # one, so P (v + 1 b_v^T) = P v + b_v and b_v moves into the output projection's bias
            "wv": wv.to(BF16).contiguous(), "wo": wo.to(BF16).contiguous(), "bo": (bo + wo @ bv).contiguous()}


def pack_weights(m: AutoencoderKL) -> dict:
    dev = m.quant_conv.weight.device
    e, d = m.encoder, m.decoder
    P: dict = {}
    P["e_in_w"], P["e_in_b"] = pack.pack_conv3x3(e.conv_in.weight, _CPAD), _f32(e.conv_in.bias)
    P["e_down"] = []
    for blk in e.down_blocks:
        ds = None
        if hasattr(blk, "downsamplers"):
            cv = blk.downsamplers[0].conv
            ds = (pack.pack_conv3x3(cv.weight), _f32(cv.bias))
        P["e_down"].append(([_pack_resnet(r) for r in blk.resnets], ds))
    P["e_mid"] = _pack_mid(e.mid_block)
    P["e_no_g"], P["e_no_b"] = _f32(e.conv_norm_out.weight), _f32(e.conv_norm_out.bias)
    # quant_conv (1x1) after conv_out (3x3) is one 3x3 convolution; only the mean half of the moments is used
    # (autoencoder.py:22-24), and the scale factor multiplies it
    wq = m.quant_conv.weight.detach().float().reshape(2 * LATENT, 2 * LATENT)[:LATENT] * AutoEncoder.scale_factor
    bq = m.quant_conv.bias.detach().float()[:LATENT] * AutoEncoder.scale_factor
    w_out = torch.einsum("om,mcyx->ocyx", wq, e.conv_out.weight.detach().float())
    P["e_out_w"] = pack.pack_conv3x3(w_out)
    P["e_out_b"] = (wq @ e.conv_out.bias.detach().float() + bq).contiguous()
    # decoder: post_quant_conv (1x1 on the latent / scale_factor) as a 64 -> 64 channel GEMM on the padded rows
    wpq = torch.zeros(_CPAD, _CPAD, device=dev)
    wpq[:LATENT, :LATENT] = m.post_quant_conv.weight.detach().float().reshape(LATENT, LATENT) / AutoEncoder.scale_factor
    bpq = torch.zeros(_CPAD, device=dev)
    bpq[:LATENT] = m.post_quant_conv.bias.detach().float()
    P["d_pq_w"], P["d_pq_b"] = wpq.to(BF16).contiguous(), bpq.contiguous()
    P["d_in_w"], P["d_in_b"] = pack.pack_conv3x3(d.conv_in.weight, _CPAD), _f32(d.conv_in.bias)
    P["d_mid"] = _pack_mid(d.mid_block)
    P["d_up"] = []
    for blk in d.up_blocks:
        us = None
        if hasattr(blk, "upsamplers"):
            cv = blk.upsamplers[0].conv
            us = (pack.pack_upconv_phases(cv.weight), _f32(cv.bias))
        P["d_up"].append(([_pack_resnet(r) for r in blk.resnets], us))
    P["d_no_g"], P["d_no_b"] = _f32(d.conv_norm_out.weight), _f32(d.conv_norm_out.bias)
    w3 = torch.zeros(4, *d.conv_out.weight.shape[1:], device=dev)      # 3 -> 4 output channels (fp32 rows of 16 bytes)
    w3[:3] = d.conv_out.weight.detach().float()
    b3 = torch.zeros(4, device=dev)
    b3[:3] = d.conv_out.bias.detach().float()
    P["d_out_w"], P["d_out_b"] = pack.pack_conv3x3(w3), b3.contiguous()
    return P


# ---- launch sequences ------------------------------------------------------------------------------------------------
class _Run:
    """One encode / decode of ``n`` images: helpers over the zero-bordered flat layout [n (H+2) (W+2), C]."""

    def __init__(self, n: int, dev):
        self.n, self.dev = n, dev

    def rows(self, H, W):
        return self.n * (H + 2) * (W + 2)

    def conv(self, x, H, W, w9, b, residual=None, out_fp32=False):
        return ops.conv3x3_padded(x, w9, b, self.n, H, W, residual=residual, out_fp32=out_fp32)

    def conv1x1(self, x, H, W, w, b):
        """1x1 convolution on padded rows; the border rows stay zero."""
        out = torch.empty((x.shape[0], w.shape[0]), device=self.dev, dtype=BF16)
        g = ops.make_gemm(A_ptr=x.data_ptr(), lda=x.stride(0), a_rows=x.shape[0], W_ptr=w.data_ptr(), ldw=w.stride(0),
                          M=x.shape[0], N=w.shape[0], Kc=x.shape[1], out_ptr=out.data_ptr(), ldo=out.stride(0),
                          bias_ptr=b.data_ptr(), mask_border=True, Hp=H + 2, Wp=W + 2)
        ops.run_gemm(g)
        return out

    def gn(self, x, H, W, g, b, silu=True, out_padded=True):
        return ops.groupnorm(x, g, b, self.n, H, W, groups=GROUPS, eps=EPS, in_padded=True, out_padded=out_padded,
                             silu=silu)

    def resnet(self, x, H, W, p):
        h = self.conv(self.gn(x, H, W, p["g1"], p["b1"]), H, W, p["w1"], p["c1"])
        skip = x if p["ws"] is None else self.conv1x1(x, H, W, p["ws"], p["cs"])
        return self.conv(self.gn(h, H, W, p["g2"], p["b2"]), H, W, p["w2"], p["c2"], residual=skip)

    def attention(self, x, H, W, p):
        """x + to_out(softmax(q k^T / sqrt(C)) v) over the H W tokens of each image, one head of width C."""
        n, hw, c = self.n, H * W, x.shape[1]
        if hw % 8:
            raise SevaB200Error(f"VAE attention needs H*W % 8 == 0 at the latent resolution (got {H}x{W})")
        t = self.gn(x, H, W, p["g"], p["b"], silu=False, out_padded=False)      # tokens [n hw, c]
        qk = ops.linear(t, p["wqk"], p["bqk"])                                  # [n hw, 2c]  (q pre-scaled | k)
        o = torch.empty((n * hw, c), device=self.dev, dtype=BF16)
        s = torch.empty((hw, hw), device=self.dev, dtype=F32)
        pr = torch.empty((hw, hw), device=self.dev, dtype=BF16)
        vt = torch.empty((c, hw), device=self.dev, dtype=BF16)
        for i in range(n):
            ti, qki = t[i * hw:(i + 1) * hw], qk[i * hw:(i + 1) * hw]
            ops.linear(p["wv"], ti, None, out=vt)                               # v^T = W_v t^T  [c, hw]
            ops.linear(qki[:, :c], qki[:, c:], None, out=s, out_fp32=True)      # scores, fp32
            ops.softmax_rows(s, out=pr)
            ops.linear(pr, vt, None, out=o[i * hw:(i + 1) * hw])                # P v
        out = torch.zeros((x.shape[0], c), device=self.dev, dtype=BF16)         # token rows -> padded rows, + x
        g = ops.make_gemm(A_ptr=o.data_ptr(), lda=o.stride(0), a_rows=o.shape[0], W_ptr=p["wo"].data_ptr(),
                          ldw=p["wo"].stride(0), M=o.shape[0], N=c, Kc=c, out_ptr=out.data_ptr(), ldo=out.stride(0),
                          bias_ptr=p["bo"].data_ptr(), residual_ptr=x.data_ptr(), ldr=x.stride(0), out_map=1, H=H, Wd=W)
        ops.run_gemm(g)
        return out

    def mid(self, x, H, W, p):
        x = self.resnet(x, H, W, p["r0"])
        x = self.attention(x, H, W, p)
        return self.resnet(x, H, W, p["r1"])

    def down(self, x, H, W, w9, b):
        """Downsample2D: pad right / bottom by one, 3x3 stride 2 (gather + GEMM, rows mapped back to the padded layout)."""
        c = x.shape[1]
        ho, wo = H // 2, W // 2
        col = torch.empty((self.n * ho * wo, 9 * c), device=self.dev, dtype=BF16)
        ops.im2col3x3_s2(x, col, self.n, H, W, c, origin=1)
        out = torch.zeros((self.rows(ho, wo), w9.shape[0]), device=self.dev, dtype=BF16)
        g = ops.make_gemm(A_ptr=col.data_ptr(), lda=col.stride(0), a_rows=col.shape[0], W_ptr=w9.data_ptr(),
                          ldw=w9.stride(0), M=col.shape[0], N=w9.shape[0], Kc=9 * c, out_ptr=out.data_ptr(),
                          ldo=out.stride(0), bias_ptr=b.data_ptr(), out_map=1, H=ho, Wd=wo)
        ops.run_gemm(g)
        return out

    def up(self, x, H, W, wph, b):
        """Upsample2D: conv3x3(nearest2x(x)) as four 2x2-tap phase convolutions on x itself (SevaGemm.out_map = 2..5)."""
        c = x.shape[1]
        out = torch.zeros((self.rows(2 * H, 2 * W), wph[0].shape[0]), device=self.dev, dtype=BF16)
        for ph in range(4):
            g = ops.make_gemm(A_ptr=x.data_ptr(), lda=x.stride(0), a_rows=x.shape[0], W_ptr=wph[ph].data_ptr(),
                              ldw=wph[ph].stride(0), M=x.shape[0], N=wph[ph].shape[0], Kc=c, out_ptr=out.data_ptr(),
                              ldo=out.stride(0), ntaps=4, tap_wp=W + 2, bias_ptr=b.data_ptr(), out_map=2 + ph, H=H, Wd=W)
            ops.run_gemm(g)
        return out


def _encode_cuda(P: dict, x: torch.Tensor) -> torch.Tensor:
    n, c, H, W = x.shape
    if c != 3 or H % 8 or W % 8:
        raise SevaB200Error(f"AutoEncoder.encode expects [N, 3, H, W] with H, W multiples of 8 (got {tuple(x.shape)})")
    r = _Run(n, x.device)
    h = torch.empty((r.rows(H, W), _CPAD), device=x.device, dtype=BF16)
    ops.pack_nchw(x.float(), h)
    h = r.conv(h, H, W, P["e_in_w"], P["e_in_b"])
    for resnets, ds in P["e_down"]:
        for p in resnets:
            h = r.resnet(h, H, W, p)
        if ds is not None:
            h = r.down(h, H, W, *ds)
            H, W = H // 2, W // 2
    h = r.mid(h, H, W, P["e_mid"])
    h = r.gn(h, H, W, P["e_no_g"], P["e_no_b"])
    z = r.conv(h, H, W, P["e_out_w"], P["e_out_b"], out_fp32=True)            # mean * scale_factor, [rows, 4] fp32
    return ops.unpack_nchw(z, n, LATENT, H, W)


def _decode_cuda(P: dict, z: torch.Tensor) -> torch.Tensor:
    n, c, H, W = z.shape
    if c != LATENT:
        raise SevaB200Error(f"AutoEncoder.decode expects [N, {LATENT}, h, w] latents (got {tuple(z.shape)})")
    r = _Run(n, z.device)
    h = torch.empty((r.rows(H, W), _CPAD), device=z.device, dtype=BF16)
    ops.pack_nchw(z.float(), h)
    h = r.conv1x1(h, H, W, P["d_pq_w"], P["d_pq_b"])                           # post_quant_conv(z / scale_factor)
    h = r.conv(h, H, W, P["d_in_w"], P["d_in_b"])
    h = r.mid(h, H, W, P["d_mid"])
    for resnets, us in P["d_up"]:
        for p in resnets:
            h = r.resnet(h, H, W, p)
        if us is not None:
            h = r.up(h, H, W, *us)
            H, W = 2 * H, 2 * W
    h = r.gn(h, H, W, P["d_no_g"], P["d_no_b"])
    img = r.conv(h, H, W, P["d_out_w"], P["d_out_b"], out_fp32=True)           # [rows, 4] fp32, channel 3 is padding
    return ops.unpack_nchw(img, n, 3, H, W)


class AutoEncoder(nn.Module):
    scale_factor: float = 0.18215   # autoencoder.py:7
    downsample: int = 8             # autoencoder.py:8

    def __init__(self, chunk_size: int | None = None, state_dict: dict | None = None, seed: int | None = 0):
        super().__init__()
        if seed is not None:
            with torch.random.fork_rng(devices=[]):
                torch.manual_seed(seed)
                self.module = AutoencoderKL()
        else:
            self.module = AutoencoderKL()
        self.module.eval().requires_grad_(False)
        self.chunk_size = chunk_size
        self._packed = None
        self.register_load_state_dict_post_hook(lambda m, k: m.invalidate())
        if state_dict is not None:
            self.load_vae_state_dict(state_dict)

    # ---- weights ---------------------------------------------------------------------------------------------------
    def load_vae_state_dict(self, sd: dict, strict: bool = True):
        """Load a diffusers ``AutoencoderKL`` state dict (current or pre-0.19 attention names, 4-D attention weights)."""
        fixed = {}
        for k, v in sd.items():
            parts = k.split(".")
            if "attentions" in parts:
                parts = [_OLD_ATTN.get(q, q) for q in parts]
                k = ".".join(parts)
                if v.ndim == 4 and k.endswith("weight") and "group_norm" not in k:
                    v = v[:, :, 0, 0]
            fixed[k] = v
        out = self.module.load_state_dict(fixed, strict=strict)
        self.invalidate()
        return out

    def invalidate(self) -> None:
        self._packed = None

    def _apply(self, fn, *a, **k):
        self.invalidate()
        return super()._apply(fn, *a, **k)

    def _weights(self) -> dict:
        dev = self.module.quant_conv.weight.device
        if dev.type != "cuda":
            raise SevaB200Error("AutoEncoder parameters are on %s: seva_b200 runs on sm_100a only (no CPU fallback); "
                                "move the module with .to('cuda')" % dev)
        if self._packed is None:
            ops.check_device()
            with torch.inference_mode(False), torch.no_grad(), torch.cuda.device(dev):
                self._packed = pack_weights(self.module)
        return self._packed

    def _check(self, t: torch.Tensor, what: str) -> None:
        if not t.is_cuda:
            raise SevaB200Error(f"AutoEncoder.{what} needs CUDA tensors (seva_b200 has no CPU fallback)")
        dev = self.module.quant_conv.weight.device
        if t.device != dev:
            raise SevaB200Error(f"AutoEncoder.{what}: input on {t.device}, parameters on {dev}")

    # ---- reference surface (autoencoder.py:21-52) ---------------------------------------------------------------------
    @torch.compiler.disable
    def _encode(self, x: torch.Tensor) -> torch.Tensor:
        self._check(x, "encode")
        with torch.autocast("cuda", enabled=False), torch.cuda.device(x.device), torch.no_grad():
            return _encode_cuda(self._weights(), x).to(x.dtype if x.dtype.is_floating_point else F32)

    def encode(self, x: torch.Tensor, chunk_size: int | None = None) -> torch.Tensor:
        chunk_size = chunk_size or self.chunk_size
        if chunk_size is not None:
            return torch.cat([self._encode(x_chunk) for x_chunk in x.split(chunk_size)], dim=0)
        return self._encode(x)

    @torch.compiler.disable
    def _decode(self, z: torch.Tensor) -> torch.Tensor:
        self._check(z, "decode")
        with torch.autocast("cuda", enabled=False), torch.cuda.device(z.device), torch.no_grad():
            return _decode_cuda(self._weights(), z).to(z.dtype if z.dtype.is_floating_point else F32)

    def decode(self, z: torch.Tensor, chunk_size: int | None = None) -> torch.Tensor:
        chunk_size = chunk_size or self.chunk_size
        if chunk_size is not None:
            return torch.cat([self._decode(z_chunk) for z_chunk in z.split(chunk_size)], dim=0)
        return self._decode(z)

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        return self.decode(self.encode(x))
