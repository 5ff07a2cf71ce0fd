"""SURVEY 8f N3, oracle side: the encoder / decoder restatement in oracle/vae_oracle.py against an INDEPENDENT implementation
of the same architecture.  diffusers (whose AutoencoderKL the reference wraps) is not in this image; the LDM-lineage
`Encoder` / `Decoder` are -- torchtitan ships them for the FLUX autoencoder (torchtitan/experiments/flux/model/autoencoder.py),
and with ch 128, ch_mult (1, 2, 4, 4), two residual blocks per level and z_channels 4 they ARE the Stable Diffusion VAE's
encoder / decoder, the code diffusers' AutoencoderKL was converted from.  The key mapping below is that conversion
(diffusers `convert_ldm_vae_checkpoint`) read backwards: down_blocks.i.resnets.j <-> down.i.block.j, conv_shortcut <->
nin_shortcut, downsamplers.0 <-> downsample, mid_block.resnets.{0,1} <-> mid.block_{1,2}, mid_block.attentions.0
{group_norm, to_q, to_k, to_v, to_out.0} <-> mid.attn_1 {norm, q, k, v, proj_out} (linear <-> 1x1 conv),
conv_norm_out <-> norm_out, and up_blocks.i <-> up.(3 - i).

Pinned by this: `vae_oracle.encoder` and `vae_oracle.decoder` (all the arithmetic of the two networks).  Not pinned: the thin
wrapper around them (quant / post-quant 1x1 convs, `latent_dist.mean`, the 0.18215 scale), which is restated from diffusers'
published source; DESIGN.md keeps "parity unpinned" for the module as a whole.
"""
import os
import sys

import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "stable-virtual-camera_b200")]

ldm = pytest.importorskip("torchtitan.experiments.flux.model.autoencoder")


def _resnet(dst: dict, d: str, sd: dict, s: str) -> None:
    for n in ("norm1", "conv1", "norm2", "conv2"):
        for wb in ("weight", "bias"):
            dst[f"{d}.{n}.{wb}"] = sd[f"{s}.{n}.{wb}"]
    if f"{s}.conv_shortcut.weight" in sd:
        dst[f"{d}.nin_shortcut.weight"] = sd[f"{s}.conv_shortcut.weight"]
        dst[f"{d}.nin_shortcut.bias"] = sd[f"{s}.conv_shortcut.bias"]


def _mid(dst: dict, sd: dict, side: str) -> None:
    _resnet(dst, "mid.block_1", sd, f"{side}.mid_block.resnets.0")
    _resnet(dst, "mid.block_2", sd, f"{side}.mid_block.resnets.1")
    a = f"{side}.mid_block.attentions.0"
    dst["mid.attn_1.norm.weight"], dst["mid.attn_1.norm.bias"] = sd[f"{a}.group_norm.weight"], sd[f"{a}.group_norm.bias"]
    for theirs, ours in (("q", "to_q"), ("k", "to_k"), ("v", "to_v"), ("proj_out", "to_out.0")):
        dst[f"mid.attn_1.{theirs}.weight"] = sd[f"{a}.{ours}.weight"][:, :, None, None]
        dst[f"mid.attn_1.{theirs}.bias"] = sd[f"{a}.{ours}.bias"]


def _encoder_keys(sd: dict, levels: int, layers: int) -> dict:
    d: dict = {}
    for n in ("conv_in", "conv_out"):
        d[f"{n}.weight"], d[f"{n}.bias"] = sd[f"encoder.{n}.weight"], sd[f"encoder.{n}.bias"]
    d["norm_out.weight"], d["norm_out.bias"] = sd["encoder.conv_norm_out.weight"], sd["encoder.conv_norm_out.bias"]
    for i in range(levels):
        for j in range(layers):
            _resnet(d, f"down.{i}.block.{j}", sd, f"encoder.down_blocks.{i}.resnets.{j}")
        if i != levels - 1:
            for wb in ("weight", "bias"):
                d[f"down.{i}.downsample.conv.{wb}"] = sd[f"encoder.down_blocks.{i}.downsamplers.0.conv.{wb}"]
    _mid(d, sd, "encoder")
    return d


def _decoder_keys(sd: dict, levels: int, layers: int) -> dict:
    d: dict = {}
    for n in ("conv_in", "conv_out"):
        d[f"{n}.weight"], d[f"{n}.bias"] = sd[f"decoder.{n}.weight"], sd[f"decoder.{n}.bias"]
    d["norm_out.weight"], d["norm_out.bias"] = sd["decoder.conv_norm_out.weight"], sd["decoder.conv_norm_out.bias"]
    for i in range(levels):          # diffusers runs up_blocks 0..3; the LDM decoder runs up.3 .. up.0
        for j in range(layers + 1):
            _resnet(d, f"up.{levels - 1 - i}.block.{j}", sd, f"decoder.up_blocks.{i}.resnets.{j}")
        if i != levels - 1:
            for wb in ("weight", "bias"):
                d[f"up.{levels - 1 - i}.upsample.conv.{wb}"] = sd[f"decoder.up_blocks.{i}.upsamplers.0.conv.{wb}"]
    _mid(d, sd, "decoder")
    return d


def _rel(a, b):
    return ((a - b).norm() / b.norm()).item()


def test_oracle_vae_encoder_decoder_match_the_ldm_implementation():
    from oracle import vae_oracle as VO

    sd = VO.random_state_dict(seed=5)
    levels, layers = len(VO.BLOCK_OUT), VO.LAYERS_PER_BLOCK
    mult = tuple(c // VO.BLOCK_OUT[0] for c in VO.BLOCK_OUT)
    enc = ldm.Encoder(resolution=64, in_channels=3, ch=VO.BLOCK_OUT[0], ch_mult=mult, num_res_blocks=layers,
                      z_channels=VO.LATENT).eval()
    dec = ldm.Decoder(ch=VO.BLOCK_OUT[0], out_ch=3, ch_mult=mult, num_res_blocks=layers, in_channels=3, resolution=64,
                      z_channels=VO.LATENT).eval()
    enc.load_state_dict(_encoder_keys(sd, levels, layers), strict=True)   # strict: every parameter of theirs is mapped
    dec.load_state_dict(_decoder_keys(sd, levels, layers), strict=True)
    g = torch.Generator().manual_seed(1)
    x = torch.randn(2, 3, 32, 24, generator=g)       # non-square, two frames
    z = torch.randn(2, VO.LATENT, 4, 3, generator=g)
    with torch.no_grad():
        e_ours, e_theirs = VO.encoder(x, sd), enc(x)
        d_ours, d_theirs = VO.decoder(z, sd), dec(z)
    assert e_ours.shape == e_theirs.shape == (2, 2 * VO.LATENT, 4, 3)
    assert d_ours.shape == d_theirs.shape == (2, 3, 32, 24)
    assert _rel(e_ours, e_theirs) < 2e-5, _rel(e_ours, e_theirs)
    assert _rel(d_ours, d_theirs) < 2e-5, _rel(d_ours, d_theirs)
    used = {k for k in sd if k.startswith(("encoder.", "decoder."))}
    assert len(_encoder_keys(sd, levels, layers)) + len(_decoder_keys(sd, levels, layers)) == len(used)  # nothing of ours left out
