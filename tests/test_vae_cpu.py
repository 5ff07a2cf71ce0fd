"""N3: the VAE either side of the sampling loop -- host side and oracle, no GPU.

diffusers (the library the reference wraps, seva/modules/autoencoder.py:2,13) is absent here, so the oracle is a
restatement with parity UNPINNED (oracle/vae_oracle.py header).  What these tests pin: the parameter tree carries
diffusers' names and the SD VAE's 83 653 863 parameters, old-format attention checkpoints load, the oracle's encode /
decode have the reference's shapes, scale factor and chunking, and the product refuses CPU tensors."""
import pytest
import torch

from oracle import vae_oracle as VO
from seva_b200._lib import SevaB200Error
from seva_b200.autoencoder import AutoEncoder, AutoencoderKL


def test_parameter_tree_has_diffusers_names_and_sd_vae_size():
    sd = AutoencoderKL().state_dict()
    shapes = VO.param_shapes()
    assert set(sd) == set(shapes)
    assert all(tuple(sd[k].shape) == tuple(shapes[k]) for k in sd)
    assert sum(v.numel() for v in sd.values()) == 83_653_863   # the published size of SD's AutoencoderKL
    for k in ("encoder.down_blocks.0.downsamplers.0.conv.weight", "decoder.up_blocks.2.upsamplers.0.conv.bias",
              "encoder.mid_block.attentions.0.to_out.0.weight", "decoder.up_blocks.2.resnets.0.conv_shortcut.weight",
              "quant_conv.weight", "post_quant_conv.bias", "decoder.conv_norm_out.weight"):
        assert k in sd
    assert "encoder.down_blocks.3.downsamplers.0.conv.weight" not in sd
    assert "decoder.up_blocks.3.upsamplers.0.conv.weight" not in sd


def test_old_attention_names_load():
    ref = VO.random_state_dict(3)
    old = {}
    ren = {"to_q": "query", "to_k": "key", "to_v": "value", "to_out.0": "proj_attn"}
    for k, v in ref.items():
        if ".attentions.0." in k and "group_norm" not in k:
            for new, o in ren.items():
                if f".{new}." in k:
                    k = k.replace(f".{new}.", f".{o}.")
                    if k.endswith("weight"):
                        v = v[:, :, None, None]     # the deprecated AttentionBlock checkpoints hold 1x1 conv weights
                    break
        old[k] = v
    ae = AutoEncoder(seed=None)
    res = ae.load_vae_state_dict(old)
    assert not res.missing_keys and not res.unexpected_keys
    got = ae.module.state_dict()
    assert all(torch.equal(got[k], ref[k]) for k in ref)


def test_oracle_shapes_scale_and_chunking():
    sd = VO.random_state_dict(1)
    g = torch.Generator().manual_seed(0)
    x = torch.rand(3, 3, 16, 24, generator=g) * 2 - 1
    with torch.no_grad():
        z = VO.encode(x, sd)
        assert z.shape == (3, 4, 2, 3)
        zc = VO.encode(x, sd, chunk_size=2)
        torch.testing.assert_close(z, zc, rtol=1e-5, atol=1e-6)
        # mean * scale_factor (autoencoder.py:21-25)
        mom = torch.nn.functional.conv2d(VO.encoder(x, sd), sd["quant_conv.weight"], sd["quant_conv.bias"])
        torch.testing.assert_close(z, mom[:, :4] * 0.18215, rtol=1e-6, atol=1e-7)
        y = VO.decode(z, sd)
        assert y.shape == (3, 3, 16, 24)
        yc = VO.decode(z, sd, chunk_size=1)
        torch.testing.assert_close(y, yc, rtol=1e-4, atol=2e-5)   # batch size changes the conv algorithm (fp32 summation order)


def test_product_refuses_cpu_tensors():
    ae = AutoEncoder(chunk_size=1)
    assert ae.scale_factor == 0.18215 and ae.downsample == 8
    with pytest.raises(SevaB200Error):
        ae.encode(torch.zeros(1, 3, 16, 16))
    with pytest.raises(SevaB200Error):
        ae.decode(torch.zeros(1, 4, 2, 2))
