"""Host-side preprocessing bookkeeping against fixtures generated from the REAL reference
(oracle/gen_golden_preprocess.py): resizing factors, shortest-side sizes -- equality demanded -- and the oracle's restatement
of the conditioner's image path against torch itself."""
import os
import sys

import numpy as np
import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "stable-virtual-camera_b200")]
from oracle import clip_oracle as CO  # noqa: E402
from seva_b200 import preprocess as PP  # noqa: E402

G = np.load(os.path.join(ROOT, "tests", "golden", "preprocess.npz"))


def test_resizing_factor_equals_reference():
    for (a, b, c, d), cover, fit in zip(G["rf_shapes"].tolist(), G["rf_cover"], G["rf_fit"]):
        assert PP.get_resizing_factor((a, b), (c, d), True) == cover, (a, b, c, d)
        assert PP.get_resizing_factor((a, b), (c, d), False) == fit, (a, b, c, d)


def test_shortest_side_equals_reference():
    for (w, h, s), ref in zip(G["wh_in"].tolist(), G["wh_out"].tolist()):
        assert list(PP.get_wh_with_fixed_shortest_side(w, h, s)) == ref
    assert PP.get_wh_with_fixed_shortest_side(5, 7, None) == (5, 7)
    assert PP.get_wh_with_fixed_shortest_side(5, 7, 0) == (5, 7)


def test_cpu_images_raise():
    from seva_b200._lib import SevaB200Error

    with pytest.raises(SevaB200Error):
        PP.transform_img_and_K(torch.rand(1, 3, 16, 16), (8, 8))
    with pytest.raises(AssertionError):
        PP.transform_img_and_K(torch.rand(1, 3, 16, 16), (8, 8), mode="tile")


def test_oracle_gaussian_and_attention_restatements():
    # the blur restated as two grouped convolutions equals a direct 2-D correlation with the outer-product kernel
    g = torch.Generator().manual_seed(0)
    x = torch.rand(2, 3, 40, 56, generator=g)
    ky, sy = CO.blur_params(40, 16)
    kx, sx = CO.blur_params(56, 16)
    assert ky % 2 == 1 and kx % 2 == 1 and ky >= 3 and kx >= 3
    k2 = torch.outer(CO.gaussian_kernel1d(ky, sy), CO.gaussian_kernel1d(kx, sx))
    xp = torch.nn.functional.pad(x, (kx // 2, kx // 2, ky // 2, ky // 2), mode="reflect")
    direct = torch.nn.functional.conv2d(xp, k2.view(1, 1, ky, kx).expand(3, 1, ky, kx), groups=3)
    direct = torch.nn.functional.interpolate(direct, size=(16, 16), mode="bicubic", align_corners=True)
    assert torch.allclose(CO.resize_bicubic_aa(x, (16, 16)), direct, atol=1e-5)
    # upscaling: no blur
    assert torch.equal(CO.resize_bicubic_aa(x, (80, 112)),
                       torch.nn.functional.interpolate(x, size=(80, 112), mode="bicubic", align_corners=True))
    # the attention restatement equals nn.MultiheadAttention (what open_clip's ResidualAttentionBlock calls)
    width, heads, L = 64, 4, 10
    mha = torch.nn.MultiheadAttention(width, heads, batch_first=True)
    h = torch.randn(2, L, width, generator=g)
    ref = mha(h, h, h, need_weights=False)[0]
    sd = {"conv1.weight": torch.zeros(width, 3, 2, 2), "class_embedding": torch.zeros(width),
          "positional_embedding": torch.zeros(5, width), "ln_pre.weight": torch.ones(width), "ln_pre.bias": torch.zeros(width),
          "ln_post.weight": torch.ones(width), "ln_post.bias": torch.zeros(width), "proj": torch.eye(width)}
    qkv = torch.nn.functional.linear(h, mha.in_proj_weight, mha.in_proj_bias)
    q, k, v = (t.reshape(2, L, heads, width // heads).transpose(1, 2) for t in qkv.chunk(3, dim=-1))
    a = torch.softmax((q @ k.transpose(-1, -2)) / (width // heads) ** 0.5, dim=-1) @ v
    got = torch.nn.functional.linear(a.transpose(1, 2).reshape(2, L, width), mha.out_proj.weight, mha.out_proj.bias)
    assert torch.allclose(got, ref, atol=1e-5)
    assert sd  # keys document the state-dict names vit_forward consumes


def test_oracle_vit_shapes():
    from seva_b200.conditioner import CLIPConditioner, ViTParams

    p = ViTParams(image_size=28, patch_size=14, width=64, layers=2, heads=2, embed_dim=32)
    c = CLIPConditioner(p, seed=3)
    sd = c.module.state_dict()
    assert "visual.transformer.resblocks.1.attn.in_proj_weight" in sd and "visual.proj" in sd
    x = torch.rand(2, 3, 40, 40, generator=torch.Generator().manual_seed(1)) * 2 - 1
    y = CO.conditioner_forward(sd, x, heads=2, size=28)
    assert y.shape == (2, 32) and torch.isfinite(y).all()
    # key set of a tower: 12 tensors per block + 8 (conv1, class / positional embedding, ln_pre, ln_post, proj)
    assert len(sd) == 2 * 12 + 8
