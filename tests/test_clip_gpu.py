"""SURVEY 8f N4 on the GPU: the CLIP conditioner's image path and ViT tower, and the input preprocessing, through the public
classes / the C ABI, against the oracle restatement (oracle/clip_oracle.py; parity unpinned: open_clip and kornia are absent
from this image) and -- for load_img_and_K / transform_img_and_K -- against fixtures generated from the real reference."""
import os
import sys

import numpy as np
import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "stable-virtual-camera_b200")]

pytestmark = pytest.mark.gpu


def _rel(a, b):
    return ((a.float() - b.float()).norm() / b.float().norm()).item()


@pytest.fixture(autouse=True)
def _no_tf32():
    old = (torch.backends.cuda.matmul.allow_tf32, torch.backends.cudnn.allow_tf32)
    torch.backends.cuda.matmul.allow_tf32 = False
    torch.backends.cudnn.allow_tf32 = False
    yield
    torch.backends.cuda.matmul.allow_tf32, torch.backends.cudnn.allow_tf32 = old


@pytest.mark.parametrize("shape,size", [((2, 3, 576, 576), (224, 224)), ((1, 3, 300, 400), (224, 224)),
                                        ((1, 3, 100, 120), (224, 224)), ((2, 3, 1024, 576), (224, 224)),
                                        ((1, 3, 230, 225), (224, 224))])
def test_resize_bicubic_antialias(shape, size):
    from oracle import clip_oracle as CO
    from seva_b200 import ops

    x = (torch.rand(shape, generator=torch.Generator().manual_seed(0)) * 2 - 1).cuda()
    got = ops.resize_bicubic_aa(x, size)
    ref = CO.resize_bicubic_aa(x, size)
    assert got.shape == ref.shape
    assert (got - ref).abs().max().item() < 2e-5, (got - ref).abs().max().item()


@pytest.mark.parametrize("shape,size", [((1, 3, 37, 53), (24, 32)), ((2, 3, 576, 1024), (324, 576)), ((1, 3, 64, 48), (64, 48)),
                                        ((1, 3, 50, 50), (80, 70))])
def test_resize_area(shape, size):
    from seva_b200 import ops

    x = torch.rand(shape, generator=torch.Generator().manual_seed(1)).cuda()
    got = ops.resize_area(x, size)
    ref = torch.nn.functional.interpolate(x, size, mode="area")
    assert (got - ref).abs().max().item() < 1e-6


def test_clip_patchify():
    from oracle import clip_oracle as CO
    from seva_b200 import ops

    x = (torch.rand((2, 3, 224, 224), generator=torch.Generator().manual_seed(2)) * 2 - 1).cuda()
    Lp, kp = 264, 640
    got = ops.clip_patchify(x, 14, Lp, kp, CO.CLIP_MEAN, CO.CLIP_STD).float().view(2, Lp, kp)
    mean = torch.tensor(CO.CLIP_MEAN, device="cuda").view(1, 3, 1, 1)
    std = torch.tensor(CO.CLIP_STD, device="cuda").view(1, 3, 1, 1)
    n = ((x + 1) / 2 - mean) / std
    ref = torch.nn.functional.unfold(n, 14, stride=14).transpose(1, 2)            # [2, 256, 588] as (c, ky, kx)
    assert (got[:, 1:257, :588] - ref.bfloat16().float()).abs().max().item() == 0.0
    assert got[:, 0].abs().max().item() == 0 and got[:, 257:].abs().max().item() == 0 and got[:, :, 588:].abs().max().item() == 0


def test_preprocess_matches_reference_fixtures():
    from oracle.gen_golden_preprocess import LOAD_CASES, TRANSFORM_CASES, make_K
    from seva_b200 import preprocess as PP

    G = np.load(os.path.join(ROOT, "tests", "golden", "preprocess.npz"))
    for i, (h, w, size, scale, center, norm, stride, mode) in enumerate(TRANSFORM_CASES):
        img = torch.from_numpy(G[f"t{i}_img"]).cuda()
        o, K = PP.transform_img_and_K(img, size, scale=scale, center=center, K=make_K(norm, w, h, True), size_stride=stride,
                                      mode=mode)
        ref = torch.from_numpy(G[f"t{i}_out"])
        assert tuple(o.shape) == tuple(ref.shape), (i, o.shape, ref.shape)
        assert (o.cpu() - ref).abs().max().item() < 1e-6, i
        assert torch.equal(K, torch.from_numpy(G[f"t{i}_K"])), i
    for i, (h, w, size, scale, center, norm, stride, cc) in enumerate(LOAD_CASES):
        o, K = PP.load_img_and_K(torch.Size([h, w]), size, scale=scale, center=center, K=make_K(norm, w, h, False),
                                 size_stride=stride, center_crop=cc, device="cuda")
        ref = torch.from_numpy(G[f"l{i}_out"])
        assert tuple(o.shape) == tuple(ref.shape), (i, o.shape, ref.shape)
        assert (o.cpu() - ref).abs().max().item() < 1e-6, i
        assert torch.equal(K, torch.from_numpy(G[f"l{i}_K"])), i


@pytest.mark.parametrize("width,heads,layers,img", [(128, 2, 2, 96), (320, 4, 3, 120)])
def test_clip_tower_small_vs_oracle(width, heads, layers, img):
    from oracle import clip_oracle as CO
    from seva_b200.conditioner import CLIPConditioner, ViTParams

    p = ViTParams(image_size=56, patch_size=14, width=width, layers=layers, heads=heads, embed_dim=64)
    c = CLIPConditioner(p, seed=7).cuda()
    x = (torch.rand((3, 3, img, img + 8), generator=torch.Generator().manual_seed(3)) * 2 - 1).cuda()
    got = c(x)
    sd = {k: v.float() for k, v in c.module.state_dict().items()}
    ref = CO.conditioner_forward(sd, x, heads=heads, size=56)
    err = _rel(got, ref)
    print(f"CLIP tower width {width} heads {heads} (head width {width // heads}) layers {layers}: rel-L2 {err:.3e}")
    assert got.shape == (3, 64) and err < 2e-2, err
    pre = c.preprocess(x)
    assert (pre - CO.preprocess(x, 56)).abs().max().item() < 1e-4
    with pytest.raises(Exception):
        c(x.cpu())


def test_clip_vit_h14_full_size_vs_oracle():
    """The released shape (ViT-H-14: width 1280, 32 layers, 16 heads of width 80, 257 tokens) on 2 input frames @576^2."""
    from oracle import clip_oracle as CO
    from seva_b200.conditioner import CLIPConditioner

    c = CLIPConditioner(seed=11).cuda()
    x = (torch.rand((2, 3, 576, 576), generator=torch.Generator().manual_seed(4)) * 2 - 1).cuda()
    got = c(x)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    got = c(x)
    e1.record()
    torch.cuda.synchronize()
    sd = {k: v.float() for k, v in c.module.state_dict().items()}
    ref = CO.conditioner_forward(sd, x, heads=16)
    err = _rel(got, ref)
    print(f"CLIP ViT-H-14, 2 frames @576^2: rel-L2 vs fp32 oracle {err:.3e}, {e0.elapsed_time(e1):.1f} ms")
    assert got.shape == (2, 1024) and torch.isfinite(got).all() and err < 3e-2, err
