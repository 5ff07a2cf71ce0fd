"""N3: VAE encode / decode on the step's kernels vs the fp32 oracle restatement (parity UNPINNED against
diffusers itself: the library is absent, see oracle/vae_oracle.py).  Random-init SD-2.1-shaped weights, the oracle runs
on the same GPU in fp32 with TF32 off.  Tolerance: rel-L2 <= 2e-2 (bf16 activations through ~30 convolutions; the
reference itself runs the VAE under fp16 autocast, seva/eval.py:1242)."""
import pytest
import torch
import torch.nn.functional as F

from oracle import vae_oracle as VO
from seva_b200 import ops
from seva_b200.autoencoder import AutoEncoder

pytestmark = pytest.mark.gpu


def rel_l2(a, b):
    a, b = a.float(), b.float()
    return ((a - b).norm() / b.norm().clamp_min(1e-12)).item()


@pytest.fixture(scope="module")
def rig():
    torch.backends.cuda.matmul.allow_tf32 = False
    torch.backends.cudnn.allow_tf32 = False
    sd = {k: v.cuda() for k, v in VO.random_state_dict(7).items()}
    ae = AutoEncoder(state_dict=sd, seed=None).cuda()
    return sd, ae


def test_softmax_rows_kernel():
    g = torch.Generator(device="cuda").manual_seed(0)
    for rows, L in ((64, 64), (200, 5184), (3, 13000)):
        x = torch.randn(rows, L, device="cuda", generator=g) * 6
        y = ops.softmax_rows(x)
        ref = torch.softmax(x, -1)
        assert (y.float() - ref).abs().max().item() <= 2.0 ** -8 * ref.max().item() + 1e-6
        assert (y.float().sum(-1) - 1).abs().max().item() < 2e-2


def test_im2col_origin_one_is_pad_right_bottom():
    g = torch.Generator(device="cuda").manual_seed(1)
    n, h, w, c = 2, 6, 8, 16
    x = torch.randn(n, c, h, w, device="cuda", generator=g)
    xp = torch.zeros(n, h + 2, w + 2, c, device="cuda")
    xp[:, 1:-1, 1:-1] = x.permute(0, 2, 3, 1)
    xb = xp.reshape(-1, c).bfloat16()
    col = torch.empty(n * (h // 2) * (w // 2), 9 * c, device="cuda", dtype=torch.bfloat16)
    ops.im2col3x3_s2(xb, col, n, h, w, c, origin=1)
    ref = F.unfold(F.pad(xb.float().reshape(n, h + 2, w + 2, c)[:, 1:-1, 1:-1].permute(0, 3, 1, 2), (0, 1, 0, 1)), 3, stride=2)
    ref = ref.reshape(n, c, 9, -1).permute(0, 3, 2, 1).reshape(-1, 9 * c)     # [(n yo xo), tap, c]
    assert torch.equal(col.float(), ref)


@pytest.mark.parametrize("n,h,w", [(2, 8, 8), (1, 16, 8)])
def test_decode_matches_oracle(rig, n, h, w):
    sd, ae = rig
    g = torch.Generator(device="cuda").manual_seed(2)
    z = torch.randn(n, 4, h, w, device="cuda", generator=g) * 0.18215 * 4
    with torch.no_grad():
        ref = VO.decode(z, sd)
    got = ae.decode(z)
    assert got.shape == ref.shape == (n, 3, 8 * h, 8 * w)
    e = rel_l2(got, ref)
    print(f"VAE decode {n}x{h}x{w}: rel-L2 {e:.3e}")
    assert e <= 2e-2
    assert torch.equal(ae.decode(z, chunk_size=1), got)


@pytest.mark.parametrize("n,H,W", [(2, 64, 64), (1, 128, 64)])
def test_encode_matches_oracle(rig, n, H, W):
    sd, ae = rig
    g = torch.Generator(device="cuda").manual_seed(3)
    x = torch.rand(n, 3, H, W, device="cuda", generator=g) * 2 - 1
    with torch.no_grad():
        ref = VO.encode(x, sd)
    got = ae.encode(x)
    assert got.shape == ref.shape == (n, 4, H // 8, W // 8)
    e = rel_l2(got, ref)
    print(f"VAE encode {n}x{H}x{W}: rel-L2 {e:.3e}")
    assert e <= 2e-2
    assert torch.equal(ae.encode(x, chunk_size=1), got)


def test_decode_bench_shape_one_frame(rig):
    """One 576 x 576 frame (72 x 72 latent, BASELINE configs[1]); eval.py:1318 decodes decoding_t = 1 frame at a time."""
    sd, ae = rig
    g = torch.Generator(device="cuda").manual_seed(4)
    z = torch.randn(1, 4, 72, 72, device="cuda", generator=g) * 0.18215 * 4
    with torch.no_grad():
        ref = VO.decode(z, sd)
    got = ae.decode(z)
    e = rel_l2(got, ref)
    print(f"VAE decode 1x72x72 -> 576x576: rel-L2 {e:.3e}")
    assert e <= 2e-2
