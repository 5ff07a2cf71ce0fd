"""ORACLE — test infrastructure only (never imported by the product path).  PARITY UNPINNED.

CPU / plain-PyTorch fp32 restatement of what ``seva/modules/autoencoder.py:1-51`` executes: diffusers'
``AutoencoderKL`` with the ``stabilityai/stable-diffusion-2-1-base`` VAE configuration

    in/out channels 3, latent_channels 4, block_out_channels (128, 256, 512, 512), layers_per_block 2,
    norm_num_groups 32 (eps 1e-6), act "silu", mid-block attention with ONE head of width 512, quant / post-quant 1x1 convs,

wrapped by ``AutoEncoder.encode`` (``latent_dist.mean * 0.18215``, autoencoder.py:21-25) and ``AutoEncoder.decode``
(``decode(z / 0.18215).sample``, autoencoder.py:38-39), both chunked over frames (autoencoder.py:27-36, 41-50).

``diffusers`` is a third-party dependency of the reference (pyproject.toml:22, no version pinned) and is NOT installed in
this image, and the SD-2.1 weights need the network.  So this file restates the library's published block structure
(``Encoder`` / ``Decoder`` / ``DownEncoderBlock2D`` / ``UpDecoderBlock2D`` / ``UNetMidBlock2D`` / ``ResnetBlock2D`` /
``Attention`` with ``AttnProcessor2_0`` / ``Downsample2D`` / ``Upsample2D``) from its public source, over a state dict
with diffusers' parameter names, and has NO golden vectors from the library itself: parity of the CUDA path against
this file is "unpinned" in the sense of the task statement.  What is pinned: `encoder` / `decoder` against an independent
implementation of the same architecture (the LDM-lineage Encoder / Decoder shipped with torchtitan, weights mapped through
the inverse of diffusers' LDM-checkpoint conversion: tests/test_vae_ldm_cpu.py, rel-L2 < 2e-5), internal consistency (decode(encode) shapes,
the nn.Module twin below agrees with the functional form) and the CUDA path vs this file on random-init weights.
"""
from __future__ import annotations

import math

import torch
import torch.nn.functional as F

SCALE_FACTOR = 0.18215          # autoencoder.py:8
DOWNSAMPLE = 8                  # autoencoder.py:9
BLOCK_OUT = (128, 256, 512, 512)
LAYERS_PER_BLOCK = 2
LATENT = 4
GROUPS = 32
EPS = 1e-6


# ---- parameter shapes (diffusers names) -------------------------------------------------------------------------
def _resnet_shapes(prefix: str, cin: int, cout: int, out: dict) -> None:
    out[f"{prefix}.norm1.weight"] = (cin,)
    out[f"{prefix}.norm1.bias"] = (cin,)
    out[f"{prefix}.conv1.weight"] = (cout, cin, 3, 3)
    out[f"{prefix}.conv1.bias"] = (cout,)
    out[f"{prefix}.norm2.weight"] = (cout,)
    out[f"{prefix}.norm2.bias"] = (cout,)
    out[f"{prefix}.conv2.weight"] = (cout, cout, 3, 3)
    out[f"{prefix}.conv2.bias"] = (cout,)
    if cin != cout:   # ResnetBlock2D: use_in_shortcut = in_channels != out_channels -> 1x1 conv_shortcut
        out[f"{prefix}.conv_shortcut.weight"] = (cout, cin, 1, 1)
        out[f"{prefix}.conv_shortcut.bias"] = (cout,)


def _mid_shapes(prefix: str, c: int, out: dict) -> None:
    _resnet_shapes(f"{prefix}.resnets.0", c, c, out)
    a = f"{prefix}.attentions.0"
    out[f"{a}.group_norm.weight"] = (c,)
    out[f"{a}.group_norm.bias"] = (c,)
    for n in ("to_q", "to_k", "to_v", "to_out.0"):
        out[f"{a}.{n}.weight"] = (c, c)
        out[f"{a}.{n}.bias"] = (c,)
    _resnet_shapes(f"{prefix}.resnets.1", c, c, out)


def param_shapes() -> dict:
    """name -> shape of every parameter of the SD-2.1 AutoencoderKL (post-conversion attention names)."""
    s: dict = {}
    # Encoder
    s["encoder.conv_in.weight"] = (BLOCK_OUT[0], 3, 3, 3)
    s["encoder.conv_in.bias"] = (BLOCK_OUT[0],)
    cin = BLOCK_OUT[0]
    for i, cout in enumerate(BLOCK_OUT):
        for j in range(LAYERS_PER_BLOCK):
            _resnet_shapes(f"encoder.down_blocks.{i}.resnets.{j}", cin if j == 0 else cout, cout, s)
        if i != len(BLOCK_OUT) - 1:
            s[f"encoder.down_blocks.{i}.downsamplers.0.conv.weight"] = (cout, cout, 3, 3)
            s[f"encoder.down_blocks.{i}.downsamplers.0.conv.bias"] = (cout,)
        cin = cout
    _mid_shapes("encoder.mid_block", BLOCK_OUT[-1], s)
    s["encoder.conv_norm_out.weight"] = (BLOCK_OUT[-1],)
    s["encoder.conv_norm_out.bias"] = (BLOCK_OUT[-1],)
    s["encoder.conv_out.weight"] = (2 * LATENT, BLOCK_OUT[-1], 3, 3)
    s["encoder.conv_out.bias"] = (2 * LATENT,)
    # Decoder
    rev = tuple(reversed(BLOCK_OUT))
    s["decoder.conv_in.weight"] = (rev[0], LATENT, 3, 3)
    s["decoder.conv_in.bias"] = (rev[0],)
    _mid_shapes("decoder.mid_block", rev[0], s)
    cin = rev[0]
    for i, cout in enumerate(rev):
        for j in range(LAYERS_PER_BLOCK + 1):
            _resnet_shapes(f"decoder.up_blocks.{i}.resnets.{j}", cin if j == 0 else cout, cout, s)
        if i != len(rev) - 1:
            s[f"decoder.up_blocks.{i}.upsamplers.0.conv.weight"] = (cout, cout, 3, 3)
            s[f"decoder.up_blocks.{i}.upsamplers.0.conv.bias"] = (cout,)
        cin = cout
    s["decoder.conv_norm_out.weight"] = (rev[-1],)
    s["decoder.conv_norm_out.bias"] = (rev[-1],)
    s["decoder.conv_out.weight"] = (3, rev[-1], 3, 3)
    s["decoder.conv_out.bias"] = (3,)
    s["quant_conv.weight"] = (2 * LATENT, 2 * LATENT, 1, 1)
    s["quant_conv.bias"] = (2 * LATENT,)
    s["post_quant_conv.weight"] = (LATENT, LATENT, 1, 1)
    s["post_quant_conv.bias"] = (LATENT,)
    return s


def random_state_dict(seed: int = 0, dtype=torch.float32) -> dict:
    """Random-init weights of the right shapes (fan-in scaled, norms near identity): there is no network for the real ones."""
    g = torch.Generator().manual_seed(seed)
    sd = {}
    for name, shape in param_shapes().items():
        if name.endswith("norm1.weight") or name.endswith("norm2.weight") or name.endswith("norm_out.weight") or \
                name.endswith("group_norm.weight"):
            t = 1.0 + 0.1 * torch.randn(shape, generator=g)
        elif name.endswith(".bias"):
            t = 0.05 * torch.randn(shape, generator=g)
        else:
            fan_in = math.prod(shape[1:])
            t = torch.randn(shape, generator=g) / math.sqrt(fan_in)
        sd[name] = t.to(dtype)
    return sd


# ---- blocks -----------------------------------------------------------------------------------------------------
def _gn(x, sd, p):
    return F.group_norm(x, GROUPS, sd[f"{p}.weight"], sd[f"{p}.bias"], EPS)


def resnet_block(x, sd, p):
    """diffusers ResnetBlock2D.forward with temb = None, dropout 0, output_scale_factor 1."""
    h = F.conv2d(F.silu(_gn(x, sd, f"{p}.norm1")), sd[f"{p}.conv1.weight"], sd[f"{p}.conv1.bias"], padding=1)
    h = F.conv2d(F.silu(_gn(h, sd, f"{p}.norm2")), sd[f"{p}.conv2.weight"], sd[f"{p}.conv2.bias"], padding=1)
    if f"{p}.conv_shortcut.weight" in sd:
        x = F.conv2d(x, sd[f"{p}.conv_shortcut.weight"], sd[f"{p}.conv_shortcut.bias"])
    return x + h


def attention_block(x, sd, p):
    """diffusers Attention (heads = 1, dim_head = C, residual_connection, group_norm) under AttnProcessor2_0."""
    b, c, hh, ww = x.shape
    t = _gn(x, sd, f"{p}.group_norm").reshape(b, c, hh * ww).transpose(1, 2)      # [b, hw, c]
    q = F.linear(t, sd[f"{p}.to_q.weight"], sd[f"{p}.to_q.bias"])
    k = F.linear(t, sd[f"{p}.to_k.weight"], sd[f"{p}.to_k.bias"])
    v = F.linear(t, sd[f"{p}.to_v.weight"], sd[f"{p}.to_v.bias"])
    a = torch.softmax(q @ k.transpose(1, 2) / math.sqrt(c), dim=-1) @ v
    a = F.linear(a, sd[f"{p}.to_out.0.weight"], sd[f"{p}.to_out.0.bias"])
    return x + a.transpose(1, 2).reshape(b, c, hh, ww)


def mid_block(x, sd, p):
    x = resnet_block(x, sd, f"{p}.resnets.0")
    x = attention_block(x, sd, f"{p}.attentions.0")
    return resnet_block(x, sd, f"{p}.resnets.1")


def encoder(x, sd):
    h = F.conv2d(x, sd["encoder.conv_in.weight"], sd["encoder.conv_in.bias"], padding=1)
    for i in range(len(BLOCK_OUT)):
        for j in range(LAYERS_PER_BLOCK):
            h = resnet_block(h, sd, f"encoder.down_blocks.{i}.resnets.{j}")
        if i != len(BLOCK_OUT) - 1:
            # Downsample2D(use_conv, padding=0): pad right/bottom by one, then 3x3 stride 2 without padding
            h = F.pad(h, (0, 1, 0, 1))
            h = F.conv2d(h, sd[f"encoder.down_blocks.{i}.downsamplers.0.conv.weight"],
                         sd[f"encoder.down_blocks.{i}.downsamplers.0.conv.bias"], stride=2)
    h = mid_block(h, sd, "encoder.mid_block")
    h = F.silu(_gn(h, sd, "encoder.conv_norm_out"))
    return F.conv2d(h, sd["encoder.conv_out.weight"], sd["encoder.conv_out.bias"], padding=1)


def decoder(z, sd):
    h = F.conv2d(z, sd["decoder.conv_in.weight"], sd["decoder.conv_in.bias"], padding=1)
    h = mid_block(h, sd, "decoder.mid_block")
    for i in range(len(BLOCK_OUT)):
        for j in range(LAYERS_PER_BLOCK + 1):
            h = resnet_block(h, sd, f"decoder.up_blocks.{i}.resnets.{j}")
        if i != len(BLOCK_OUT) - 1:
            h = F.interpolate(h, scale_factor=2.0, mode="nearest")       # Upsample2D
            h = F.conv2d(h, sd[f"decoder.up_blocks.{i}.upsamplers.0.conv.weight"],
                         sd[f"decoder.up_blocks.{i}.upsamplers.0.conv.bias"], padding=1)
    h = F.silu(_gn(h, sd, "decoder.conv_norm_out"))
    return F.conv2d(h, sd["decoder.conv_out.weight"], sd["decoder.conv_out.bias"], padding=1)


# ---- AutoEncoder wrapper (autoencoder.py:21-51) --------------------------------------------------------------------
def _encode(x, sd):
    moments = F.conv2d(encoder(x, sd), sd["quant_conv.weight"], sd["quant_conv.bias"])
    mean, _logvar = moments.chunk(2, dim=1)          # DiagonalGaussianDistribution.mean
    return mean * SCALE_FACTOR


def _decode(z, sd):
    z = F.conv2d(z / SCALE_FACTOR, sd["post_quant_conv.weight"], sd["post_quant_conv.bias"])
    return decoder(z, sd)


def encode(x, sd, chunk_size=None):
    if chunk_size is not None:
        return torch.cat([_encode(c, sd) for c in x.split(chunk_size)], dim=0)
    return _encode(x, sd)


def decode(z, sd, chunk_size=None):
    if chunk_size is not None:
        return torch.cat([_decode(c, sd) for c in z.split(chunk_size)], dim=0)
    return _decode(z, sd)
