"""Kernel-level parity on a B200: every C-ABI kernel against a plain PyTorch fp32 statement of the same op.

Tolerances: bf16 outputs carry 2^-9 relative rounding; the gate is rel-L2 <= 5e-3 for tensor-core ops
(fp32 accumulation) and <= 4e-3 for norm/elementwise ops; fp32 kernels <= 1e-5; integer outputs bit-exact.
"""
import math

import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu


def rel_l2(a, b):
    a, b = a.float(), b.float()
    return ((a - b).norm() / b.norm().clamp_min(1e-12)).item()


@pytest.fixture(scope="module")
def ops():
    from seva_b200 import ops as o

    o.check_device()
    return o


def bf(x):
    return x.to(torch.bfloat16)


# ---------------------------------------------------------------------------------------------------
@pytest.mark.parametrize("M,K,N", [(128, 64, 32), (300, 320, 320), (1000, 1280, 640), (4097, 640, 960),
                                   (42, 1024, 1280), (2500, 320, 2560)])
def test_gemm_plain(ops, M, K, N):
    g = torch.Generator(device="cuda").manual_seed(M + K + N)
    A = bf(torch.randn(M, K, device="cuda", generator=g))
    W = bf(torch.randn(N, K, device="cuda", generator=g) / math.sqrt(K))
    b = torch.randn(N, device="cuda", generator=g)
    out = ops.linear(A, W, b)
    ref = A.float() @ W.float().t() + b
    assert rel_l2(out, ref) < 5e-3, rel_l2(out, ref)


def test_gemm_strided_residual_rowbias_fp32out(ops):
    g = torch.Generator(device="cuda").manual_seed(7)
    M, K, N, rpi = 1111, 640, 320, 101
    Abuf = bf(torch.randn(M, K + 64, device="cuda", generator=g))
    A = Abuf[:, 64:]  # column-offset view with ld = K+64
    W = bf(torch.randn(N, K, device="cuda", generator=g) / math.sqrt(K))
    b = torch.randn(N, device="cuda", generator=g)
    res_buf = bf(torch.randn(M, N + 32, device="cuda", generator=g))
    res = res_buf[:, 32:]
    groups = (M + rpi - 1) // rpi
    rb = torch.randn(groups, N, device="cuda", generator=g)
    outbuf = torch.zeros(M, N + 8, device="cuda", dtype=torch.bfloat16)
    out = ops.linear(A, W, b, out=outbuf[:, 8:], residual=res, rowbias=rb, rowbias_rpi=rpi)
    rows = torch.arange(M, device="cuda") // rpi
    ref = A.float() @ W.float().t() + b + res.float() + rb[rows]
    assert rel_l2(out, ref) < 5e-3
    assert outbuf[:, :8].abs().max().item() == 0
    out32 = ops.linear(A, W, b, out_fp32=True)
    assert rel_l2(out32, A.float() @ W.float().t() + b) < 1e-4


@pytest.mark.parametrize("M,N,reverse", [(19021, 320, False), (19021, 320, True), (8000, 960, False), (30001, 640, True)])
def test_gemm_weight_stationary_walk(ops, M, N, reverse):
    """Tall K = 320 problems take the weight-stationary walk (the CTA's N tile of W loaded once, grid a multiple of the
    N-tile count): bias + per-image bias + residual epilogue, ragged last M tile, both traversal directions."""
    g = torch.Generator(device="cuda").manual_seed(M + N)
    K, rpi = 320, 997
    A = bf(torch.randn(M, K, device="cuda", generator=g))
    W = bf(torch.randn(N, K, device="cuda", generator=g) / math.sqrt(K))
    b = torch.randn(N, device="cuda", generator=g)
    res = bf(torch.randn(M, N, device="cuda", generator=g))
    rb = torch.randn((M + rpi - 1) // rpi, N, device="cuda", generator=g)
    out = ops.linear(A, W, b, residual=res, rowbias=rb, rowbias_rpi=rpi, reverse=reverse)
    ref = A.float() @ W.float().t() + b + res.float() + rb[torch.arange(M, device="cuda") // rpi]
    assert rel_l2(out, ref) < 5e-3, rel_l2(out, ref)
    plain = ops.linear(A, W, None)
    assert rel_l2(plain, A.float() @ W.float().t()) < 5e-3


@pytest.mark.parametrize("M,K,inner", [(777, 320, 1280), (300, 640, 2560)])
def test_gemm_geglu(ops, M, K, inner):
    from seva_b200 import pack

    g = torch.Generator(device="cuda").manual_seed(11)
    A = bf(torch.randn(M, K, device="cuda", generator=g))
    W = bf(torch.randn(2 * inner, K, device="cuda", generator=g) / math.sqrt(K))
    b = torch.randn(2 * inner, device="cuda", generator=g) * 0.1
    bn = ops.pick_bn(2 * inner, True)
    Wp, bp = pack.pack_geglu(W, b, bn)
    out = ops.linear(A, Wp, bp, geglu=True)
    y = A.float() @ W.float().t() + b
    x, gate = y.chunk(2, dim=-1)
    ref = x * F.gelu(gate)
    assert out.shape == (M, inner)
    assert rel_l2(out, ref) < 5e-3, rel_l2(out, ref)


@pytest.mark.parametrize("cluster", [1, 2])
@pytest.mark.parametrize("M,K,N", [(130, 64, 32), (257, 320, 320), (1000, 640, 960), (4097, 1280, 256), (77, 128, 2560)])
def test_gemm_cluster_modes(ops, cluster, M, K, N):
    """Both tile shapes of seva_gemm_bf16 (one CTA per 128 rows / CTA pair per 256 rows) with bias, row-group bias
    and residual through the coalesced epilogue, ragged M and strided output."""
    g = torch.Generator(device="cuda").manual_seed(M * 3 + K + N + cluster)
    A = bf(torch.randn(M, K, device="cuda", generator=g))
    W = bf(torch.randn(N, K, device="cuda", generator=g) / math.sqrt(K))
    b = torch.randn(N, device="cuda", generator=g)
    res = bf(torch.randn(M, N, device="cuda", generator=g))
    rpi = 61
    rb = torch.randn((M + rpi - 1) // rpi, N, device="cuda", generator=g)
    outbuf = torch.full((M + 3, N + 16), 7.0, device="cuda", dtype=torch.bfloat16)
    out = ops.linear(A, W, b, out=outbuf[:M, 16:], residual=res, rowbias=rb, rowbias_rpi=rpi, cluster=cluster)
    rows = torch.arange(M, device="cuda") // rpi
    ref = A.float() @ W.float().t() + b + res.float() + rb[rows]
    assert rel_l2(out, ref) < 5e-3, rel_l2(out, ref)
    assert (outbuf[:, :16] == 7).all() and (outbuf[M:] == 7).all()  # nothing written outside the view
    out2 = ops.linear(A, W, None, cluster=cluster)
    assert rel_l2(out2, A.float() @ W.float().t()) < 5e-3


@pytest.mark.parametrize("cluster", [1, 2])
def test_gemm_geglu_and_conv_cluster_modes(ops, cluster):
    from seva_b200 import pack

    g = torch.Generator(device="cuda").manual_seed(100 + cluster)
    M, K, inner = 700, 320, 1280
    A = bf(torch.randn(M, K, device="cuda", generator=g))
    W = bf(torch.randn(2 * inner, K, device="cuda", generator=g) / math.sqrt(K))
    b = torch.randn(2 * inner, device="cuda", generator=g) * 0.1
    Wp, bp = pack.pack_geglu(W, b, ops.pick_bn(2 * inner, True))
    out = ops.linear(A, Wp, bp, geglu=True, cluster=cluster)
    x, gate = (A.float() @ W.float().t() + b).chunk(2, dim=-1)
    assert rel_l2(out, x * F.gelu(gate)) < 5e-3
    n_img, H, Wd, Cin, Cout = 3, 18, 10, 128, 320
    x = bf(torch.randn(n_img, Cin, H, Wd, device="cuda", generator=g))
    w = torch.randn(Cout, Cin, 3, 3, device="cuda", generator=g) / math.sqrt(9 * Cin)
    bb = torch.randn(Cout, device="cuda", generator=g)
    xp = F.pad(x.permute(0, 2, 3, 1), (0, 0, 1, 1, 1, 1)).reshape(-1, Cin).contiguous()
    res = F.pad(bf(torch.randn(n_img, H, Wd, Cout, device="cuda", generator=g)), (0, 0, 1, 1, 1, 1)).reshape(-1, Cout).contiguous()
    rb = torch.randn(n_img, Cout, device="cuda", generator=g)
    o = ops.conv3x3_padded(xp, pack.pack_conv3x3(w), bb, n_img, H, Wd, rowbias=rb, residual=res, cluster=cluster)
    o = o.reshape(n_img, H + 2, Wd + 2, Cout)
    ref = F.conv2d(x.float(), bf(w).float(), bb, padding=1) + rb[:, :, None, None]
    ref = ref.permute(0, 2, 3, 1) + res.reshape(n_img, H + 2, Wd + 2, Cout)[:, 1:-1, 1:-1].float()
    assert o[:, 0].abs().max() == 0 and o[:, -1].abs().max() == 0 and o[:, :, 0].abs().max() == 0 and o[:, :, -1].abs().max() == 0
    assert rel_l2(o[:, 1:-1, 1:-1], ref) < 5e-3


@pytest.mark.parametrize("n_img,H,Wd,Cin,Cout", [(2, 9, 9, 64, 64), (3, 18, 18, 128, 320), (2, 36, 20, 320, 640),
                                                 (1, 72, 72, 64, 320)])
def test_conv3x3(ops, n_img, H, Wd, Cin, Cout):
    from seva_b200 import pack

    g = torch.Generator(device="cuda").manual_seed(H * Cin)
    x = bf(torch.randn(n_img, Cin, H, Wd, device="cuda", generator=g))
    w = torch.randn(Cout, Cin, 3, 3, device="cuda", generator=g) / math.sqrt(9 * Cin)
    b = torch.randn(Cout, device="cuda", generator=g)
    xp = F.pad(x.permute(0, 2, 3, 1), (0, 0, 1, 1, 1, 1)).reshape(-1, Cin).contiguous()
    W9 = pack.pack_conv3x3(w)
    rb = torch.randn(n_img, Cout, device="cuda", generator=g)
    out = ops.conv3x3_padded(xp, W9, b, n_img, H, Wd, rowbias=rb)
    ref = F.conv2d(x.float(), bf(w).float(), b, padding=1) + rb[:, :, None, None]
    o = out.reshape(n_img, H + 2, Wd + 2, Cout)
    assert o[:, 0].abs().max() == 0 and o[:, -1].abs().max() == 0
    assert o[:, :, 0].abs().max() == 0 and o[:, :, -1].abs().max() == 0
    got = o[:, 1:-1, 1:-1].permute(0, 3, 1, 2)
    assert rel_l2(got, ref) < 5e-3, rel_l2(got, ref)


def test_conv3x3_small_cout_fp32(ops):
    from seva_b200 import pack

    g = torch.Generator(device="cuda").manual_seed(5)
    n_img, H, Wd, Cin, Cout = 2, 16, 24, 320, 4
    x = bf(torch.randn(n_img, Cin, H, Wd, device="cuda", generator=g))
    w = torch.randn(Cout, Cin, 3, 3, device="cuda", generator=g) / math.sqrt(9 * Cin)
    b = torch.randn(Cout, device="cuda", generator=g)
    xp = F.pad(x.permute(0, 2, 3, 1), (0, 0, 1, 1, 1, 1)).reshape(-1, Cin).contiguous()
    out = ops.conv3x3_padded(xp, pack.pack_conv3x3(w), b, n_img, H, Wd, out_fp32=True)
    ref = F.conv2d(x.float(), bf(w).float(), b, padding=1)
    got = out.reshape(n_img, H + 2, Wd + 2, Cout)[:, 1:-1, 1:-1].permute(0, 3, 1, 2)
    assert rel_l2(got, ref) < 1e-4


# ---------------------------------------------------------------------------------------------------
@pytest.mark.parametrize("B,heads,L", [(1, 1, 128), (2, 5, 300), (1, 2, 1000), (3, 1, 81), (1, 10, 2741)])
def test_attention_self(ops, B, heads, L):
    g = torch.Generator(device="cuda").manual_seed(L)
    Cc = heads * 64
    qkv = bf(torch.randn(B * L, 3 * Cc, device="cuda", generator=g))
    out = ops.attention_qkv(qkv, B, L, heads)
    q, k, v = (t.reshape(B, L, heads, 64).permute(0, 2, 1, 3).float() for t in qkv.chunk(3, dim=-1))
    ref = F.scaled_dot_product_attention(q, k, v).permute(0, 2, 1, 3).reshape(B * L, Cc)
    assert rel_l2(out, ref) < 5e-3, rel_l2(out, ref)


def test_attention_large_scores(ops):
    # peaked softmax: exercises the running-max rescale path
    g = torch.Generator(device="cuda").manual_seed(3)
    B, heads, L = 1, 2, 700
    Cc = heads * 64
    qkv = bf(torch.randn(B * L, 3 * Cc, device="cuda", generator=g) * 4.0)
    out = ops.attention_qkv(qkv, B, L, heads)
    q, k, v = (t.reshape(B, L, heads, 64).permute(0, 2, 1, 3).float() for t in qkv.chunk(3, dim=-1))
    ref = F.scaled_dot_product_attention(q, k, v).permute(0, 2, 1, 3).reshape(B * L, Cc)
    assert rel_l2(out, ref) < 8e-3, rel_l2(out, ref)


@pytest.mark.parametrize("B,heads,Lq,Lk", [(2, 5, 333, 77), (1, 2, 64, 1), (2, 1, 500, 257)])
def test_attention_cross(ops, B, heads, Lq, Lk):
    g = torch.Generator(device="cuda").manual_seed(Lq + Lk)
    Cc = heads * 64
    q = bf(torch.randn(B, Lq, Cc, device="cuda", generator=g))
    k = bf(torch.randn(B, Lk, Cc, device="cuda", generator=g))
    v = bf(torch.randn(B, Lk, Cc, device="cuda", generator=g))
    out = ops.attention(q, k, v, heads)
    sp = lambda t: t.reshape(B, -1, heads, 64).permute(0, 2, 1, 3).float()
    ref = F.scaled_dot_product_attention(sp(q), sp(k), sp(v)).permute(0, 2, 1, 3).reshape(B, Lq, Cc)
    assert rel_l2(out, ref) < 5e-3, rel_l2(out, ref)


@pytest.mark.parametrize("B,T,S,heads", [(2, 21, 81, 5), (1, 8, 324, 2), (2, 3, 50, 1), (1, 40, 30, 2), (1, 64, 20, 3),
                                         (2, 33, 17, 1), (1, 32, 9, 2), (1, 70, 6, 1)])
def test_timemix_attention(ops, B, T, S, heads):
    g = torch.Generator(device="cuda").manual_seed(T * S)
    Cc = heads * 64
    qkv = bf(torch.randn(B * T * S, 3 * Cc, device="cuda", generator=g))
    out = ops.timemix_attention(qkv, B, T, S, heads)
    # reference: "(b t) s c -> (b s) t c", SDPA over t, back
    x = qkv.reshape(B, T, S, 3 * Cc).permute(0, 2, 1, 3).reshape(B * S, T, 3 * Cc).float()
    q, k, v = (t.reshape(B * S, T, heads, 64).permute(0, 2, 1, 3) for t in x.chunk(3, dim=-1))
    o = F.scaled_dot_product_attention(q, k, v).permute(0, 2, 1, 3).reshape(B, S, T, Cc)
    ref = o.permute(0, 2, 1, 3).reshape(B * T * S, Cc)
    assert rel_l2(out, ref) < 4e-3, rel_l2(out, ref)


@pytest.mark.parametrize("B,Tq,t0,Tk,S,heads", [(2, 11, 0, 21, 81, 5), (2, 10, 11, 21, 40, 2), (1, 8, 56, 64, 30, 2),
                                                 (2, 3, 18, 21, 50, 1)])
def test_timemix_attention_kv(ops, B, Tq, t0, Tk, S, heads):
    """Frame sequence-parallel form: Tq local query frames (global frames t0..t0+Tq) against all Tk frames."""
    g = torch.Generator(device="cuda").manual_seed(Tq * S)
    Cc = heads * 64
    qkv = bf(torch.randn(B, Tk, S, 3 * Cc, device="cuda", generator=g))
    q = qkv[:, t0:t0 + Tq, :, :Cc].reshape(B * Tq * S, Cc).contiguous()
    kv = qkv[..., Cc:].reshape(B * Tk * S, 2 * Cc).contiguous()
    out = ops.timemix_attention_kv(q, kv, B, Tq, Tk, S, heads)
    full = ops.timemix_attention(qkv.reshape(B * Tk * S, 3 * Cc), B, Tk, S, heads).reshape(B, Tk, S, Cc)
    ref = full[:, t0:t0 + Tq].reshape(B * Tq * S, Cc)
    assert rel_l2(out, ref) < 1e-6 if Tk <= 32 else rel_l2(out, ref) < 4e-3, rel_l2(out, ref)


# ---------------------------------------------------------------------------------------------------
@pytest.mark.parametrize("n_img,H,Wd,Cc,in_pad,out_pad,silu,dense", [
    (2, 9, 9, 64, True, True, True, True),
    (3, 18, 18, 320, True, True, True, True),
    (2, 36, 36, 960, True, True, True, True),
    (2, 12, 10, 1280, True, False, False, False),
    (2, 8, 8, 2560, True, True, True, True),
    (1, 24, 24, 640, False, True, True, False),
])
def test_groupnorm(ops, n_img, H, Wd, Cc, in_pad, out_pad, silu, dense):
    g = torch.Generator(device="cuda").manual_seed(Cc + H)
    x = bf(torch.randn(n_img, Cc, H, Wd, device="cuda", generator=g) * 1.5 + 0.3)
    gamma = torch.randn(Cc, device="cuda", generator=g)
    beta = torch.randn(Cc, device="cuda", generator=g)
    eps = 1e-5 if silu else 1e-6
    xc = x.permute(0, 2, 3, 1)
    xin = (F.pad(xc, (0, 0, 1, 1, 1, 1)) if in_pad else xc).reshape(-1, Cc).contiguous()
    dmap = dw = db = None
    if dense:
        dmap = torch.randn(n_img, H * Wd, 6, device="cuda", generator=g)
        dw = torch.randn(2 * Cc, 6, device="cuda", generator=g) * 0.3
        db = torch.randn(2 * Cc, device="cuda", generator=g) * 0.3
    out = ops.groupnorm(xin, gamma, beta, n_img, H, Wd, eps=eps, in_padded=in_pad, out_padded=out_pad, silu=silu,
                        dense=dmap, dense_w=dw, dense_b=db)
    ref = F.group_norm(x.float(), 32, gamma, beta, eps)
    if silu:
        ref = F.silu(ref)
    if dense:
        # the 1x1 modulation conv runs on bf16 operands with fp32 accumulation, as the reference's Conv2d does under
        # torch.autocast (layers.py:124-128); the comparison rounds the operands the same way
        d = (torch.einsum("npk,ck->ncp", bf(dmap).float(), bf(dw).float()).reshape(n_img, 2 * Cc, H, Wd)
             + bf(db).float()[None, :, None, None])
        sc, sh = d.chunk(2, dim=1)
        ref = ref * (1 + sc) + sh
    if out_pad:
        o = out.reshape(n_img, H + 2, Wd + 2, Cc)
        assert o[:, 0].abs().max() == 0 and o[:, -1].abs().max() == 0
        assert o[:, :, 0].abs().max() == 0 and o[:, :, -1].abs().max() == 0
        got = o[:, 1:-1, 1:-1].permute(0, 3, 1, 2)
    else:
        got = out.reshape(n_img, H, Wd, Cc).permute(0, 3, 1, 2)
    assert rel_l2(got, ref) < 4e-3, rel_l2(got, ref)


@pytest.mark.parametrize("M,Cc", [(1000, 320), (333, 640), (50, 1280), (77, 64), (40, 192)])
def test_layernorm(ops, M, Cc):
    g = torch.Generator(device="cuda").manual_seed(Cc)
    x = bf(torch.randn(M, Cc, device="cuda", generator=g) * 2 + 0.5)
    gamma = torch.randn(Cc, device="cuda", generator=g)
    beta = torch.randn(Cc, device="cuda", generator=g)
    out = ops.layernorm(x, gamma, beta)
    ref = F.layer_norm(x.float(), (Cc,), gamma, beta, 1e-5)
    assert rel_l2(out, ref) < 4e-3


# ---------------------------------------------------------------------------------------------------
def test_resize_bilinear(ops):
    g = torch.Generator(device="cuda").manual_seed(1)
    src = torch.randn(3, 6, 72, 72, device="cuda", generator=g)
    for hd in (72, 36, 18, 9):
        out = ops.resize_bilinear_ac(src, hd, hd)
        ref = F.interpolate(src, size=(hd, hd), mode="bilinear", align_corners=True)
        ref = ref.permute(0, 2, 3, 1).reshape(3, hd * hd, 6)
        assert (out - ref).abs().max().item() < 2e-5


def test_upsample_and_im2col(ops):
    from seva_b200 import pack

    g = torch.Generator(device="cuda").manual_seed(2)
    N, h, w, Cc = 2, 9, 12, 64
    x = bf(torch.randn(N, Cc, h, w, device="cuda", generator=g))
    xp = F.pad(x.permute(0, 2, 3, 1), (0, 0, 1, 1, 1, 1)).reshape(-1, Cc).contiguous()
    up = torch.full((N * (2 * h + 2) * (2 * w + 2), Cc), 7.0, device="cuda", dtype=torch.bfloat16)
    ops.upsample2x_pad(xp, up, N, h, w, Cc)
    ref = F.pad(F.interpolate(x.float(), scale_factor=2, mode="nearest").permute(0, 2, 3, 1), (0, 0, 1, 1, 1, 1))
    assert torch.equal(up.float().reshape(ref.shape), ref)
    # stride-2 conv through im2col + GEMM
    N, h, w, Cc, Cout = 2, 12, 16, 64, 128
    x = bf(torch.randn(N, Cc, h, w, device="cuda", generator=g))
    xp = F.pad(x.permute(0, 2, 3, 1), (0, 0, 1, 1, 1, 1)).reshape(-1, Cc).contiguous()
    col = torch.empty(N * (h // 2) * (w // 2), 9 * Cc, device="cuda", dtype=torch.bfloat16)
    ops.im2col3x3_s2(xp, col, N, h, w, Cc)
    wt = torch.randn(Cout, Cc, 3, 3, device="cuda", generator=g) / math.sqrt(9 * Cc)
    b = torch.randn(Cout, device="cuda", generator=g)
    out = ops.linear(col, pack.pack_conv3x3(wt), b)
    ref = F.conv2d(x.float(), bf(wt).float(), b, stride=2, padding=1)
    got = out.reshape(N, h // 2, w // 2, Cout).permute(0, 3, 1, 2)
    assert rel_l2(got, ref) < 5e-3


@pytest.mark.parametrize("N,h,w,Cin,Cout", [(2, 9, 12, 64, 128), (3, 18, 18, 128, 96), (1, 5, 7, 192, 64)])
def test_upsample_conv_as_four_phase_convs(ops, N, h, w, Cin, Cout):
    """Upsample.forward (layers.py:42-46): conv3x3(nearest2x(x)) as four 2x2-tap phase GEMMs on the source image
    (SevaGemm.ntaps = 4, out_map = 2 + phase) vs the upsample-then-convolve statement in fp32."""
    from seva_b200 import pack
    from seva_b200.ops import make_gemm, run_gemm

    g = torch.Generator(device="cuda").manual_seed(12)
    x = bf(torch.randn(N, Cin, h, w, device="cuda", generator=g))
    xp = F.pad(x.permute(0, 2, 3, 1), (0, 0, 1, 1, 1, 1)).reshape(-1, Cin).contiguous()
    wt = torch.randn(Cout, Cin, 3, 3, device="cuda", generator=g) / math.sqrt(9 * Cin)
    b = torch.randn(Cout, device="cuda", generator=g)
    out = torch.zeros(N * (2 * h + 2) * (2 * w + 2), Cout, device="cuda", dtype=torch.bfloat16)
    phases = pack.pack_upconv_phases(wt)
    for ph in range(4):
        gd = make_gemm(A_ptr=xp.data_ptr(), lda=Cin, a_rows=xp.shape[0], W_ptr=phases[ph].data_ptr(), ldw=4 * Cin,
                       M=xp.shape[0], N=Cout, Kc=Cin, out_ptr=out.data_ptr(), ldo=Cout, ntaps=4, tap_wp=w + 2,
                       bias_ptr=b.data_ptr(), out_map=2 + ph, H=h, Wd=w)
        run_gemm(gd)
    ref = F.conv2d(F.interpolate(x.float(), scale_factor=2, mode="nearest"), wt, b, padding=1)
    o = out.reshape(N, 2 * h + 2, 2 * w + 2, Cout)
    assert rel_l2(o[:, 1:-1, 1:-1].permute(0, 3, 1, 2), ref) < 6e-3
    assert o[:, 0].abs().max() == 0 and o[:, -1].abs().max() == 0 and o[:, :, 0].abs().max() == 0 and o[:, :, -1].abs().max() == 0


def test_gemm_out_map_compact_to_padded(ops):
    g = torch.Generator(device="cuda").manual_seed(9)
    n_img, H, Wd, K, N = 2, 6, 10, 64, 64
    A = bf(torch.randn(n_img * H * Wd, K, device="cuda", generator=g))
    W = bf(torch.randn(N, K, device="cuda", generator=g) / 8)
    res = bf(torch.randn(n_img, H + 2, Wd + 2, N, device="cuda", generator=g))
    out = torch.zeros(n_img * (H + 2) * (Wd + 2), N, device="cuda", dtype=torch.bfloat16)
    from seva_b200.ops import make_gemm, run_gemm

    gd = make_gemm(A_ptr=A.data_ptr(), lda=K, a_rows=A.shape[0], W_ptr=W.data_ptr(), ldw=K, M=A.shape[0], N=N, Kc=K,
                   out_ptr=out.data_ptr(), ldo=N, residual_ptr=res.data_ptr(), ldr=N, out_map=1, H=H, Wd=Wd)
    run_gemm(gd)
    ref = (A.float() @ W.float().t()).reshape(n_img, H, Wd, N) + res[:, 1:-1, 1:-1].float()
    o = out.reshape(n_img, H + 2, Wd + 2, N)
    assert rel_l2(o[:, 1:-1, 1:-1], ref) < 5e-3
    assert o[:, 0].abs().max() == 0 and o[:, :, 0].abs().max() == 0


def test_small_dense(ops):
    g = torch.Generator(device="cuda").manual_seed(4)
    t = torch.tensor([999, 979, 19, 0, 500], device="cuda", dtype=torch.int64)
    emb = ops.timestep_embedding(t, 320)
    half = 160
    freqs = torch.exp(-math.log(10000) * torch.arange(half, dtype=torch.float32, device="cuda") / half)
    args = t[:, None].float() * freqs[None]
    ref = torch.cat([torch.cos(args), torch.sin(args)], dim=-1)
    assert (emb - ref).abs().max().item() < 2e-3  # sin/cos of arguments up to 999 rad
    x = torch.randn(42, 1280, device="cuda", generator=g)
    W = bf(torch.randn(640, 1280, device="cuda", generator=g) / 36)
    b = torch.randn(640, device="cuda", generator=g)
    out = ops.gemv_rows(x, W, b, silu_in=True)
    ref = F.silu(x) @ W.float().t() + b
    assert rel_l2(out, ref) < 1e-5
    out = ops.gemv_rows(x, W, b, silu_out=True)
    assert rel_l2(out, F.silu(x @ W.float().t() + b)) < 1e-5


def test_sampler_elementwise(ops):
    g = torch.Generator(device="cuda").manual_seed(6)
    T, h, w = 5, 16, 12
    N = 2 * T
    x = torch.randn(T, 4, h, w, device="cuda", generator=g) * 80
    replace = torch.randn(N, 5, h, w, device="cuda", generator=g)
    replace[:, 4] = (torch.rand(N, 1, 1, device="cuda", generator=g) > 0.5).float().expand(N, h, w)
    concat = torch.randn(N, 7, h, w, device="cuda", generator=g)
    table = torch.linspace(0.02, 85, 1000, device="cuda") ** 1.0
    sigma = torch.full((N,), 53.7, device="cuda")
    idx, sq, c_in, c_out = ops.sigma_quantize(sigma, table)
    ref_idx = (sigma - table[:, None]).abs().argmin(dim=0)
    assert torch.equal(idx, ref_idx)
    assert torch.equal(sq, table[ref_idx])
    assert torch.allclose(c_in, 1 / (sq ** 2 + 1.0) ** 0.5, rtol=1e-6)
    net_in = torch.full((N * (h + 2) * (w + 2), 64), 3.0, device="cuda", dtype=torch.bfloat16)
    blended = torch.empty(N, 4, h, w, device="cuda")
    ops.denoise_prepare(x, replace, concat, c_in, net_in, blended, N, T, h, w)
    xx = torch.cat([x] * 2)
    m = replace[:, 4:5]
    ref_bl = xx * (1 - m) + replace[:, :4] * m
    assert torch.allclose(blended, ref_bl, rtol=1e-6, atol=1e-6)
    ref_in = torch.cat([ref_bl * c_in[:, None, None, None], concat], 1)
    ni = net_in.reshape(N, h + 2, w + 2, 64)
    assert ni[:, 0].abs().max() == 0 and ni[:, :, -1].abs().max() == 0 and ni[..., 11:].abs().max() == 0
    assert rel_l2(ni[:, 1:-1, 1:-1, :11].permute(0, 3, 1, 2), ref_in) < 3e-3
    eps = torch.randn(N * (h + 2) * (w + 2), 4, device="cuda", generator=g)
    den = torch.empty(N, 4, h, w, device="cuda")
    ops.denoise_finish(eps, blended, c_out, den, N, h, w)
    e = eps.reshape(N, h + 2, w + 2, 4)[:, 1:-1, 1:-1].permute(0, 3, 1, 2)
    ref_den = e * c_out[:, None, None, None] + ref_bl
    assert torch.allclose(den, ref_den, rtol=1e-5, atol=1e-5)
    scale = torch.rand(T, device="cuda", generator=g) + 1.2
    sh = torch.full((T,), 53.7 + 1e-6, device="cuda")
    sn = torch.full((T,), 48.3, device="cuda")
    xn = ops.cfg_euler_step(x, den, scale, sh, sn)
    u, c = ref_den.chunk(2)
    d = u + scale[:, None, None, None] * (c - u)
    ref_x = x + (sn - sh)[:, None, None, None] * ((x - d) / sh[:, None, None, None])
    assert torch.allclose(xn, ref_x, rtol=1e-5, atol=1e-4)
    noise = torch.randn_like(x)
    xa = ops.add_noise(x, noise, sn, sh)
    assert torch.allclose(xa, x + noise * ((sh ** 2 - sn ** 2) ** 0.5)[:, None, None, None], rtol=1e-5, atol=1e-5)
    # sigma_hat == sigma in fp32 for large sigma (sigma + 1e-6 rounds back): the injected noise must be exactly 0
    from oracle import seva_oracle as O
    sig50 = O.ddpm_sigmas(50).cuda()
    for s in sig50[:10]:
        sg = torch.full((T,), float(s), device="cuda")
        xa = ops.add_noise(x, noise, sg, sg + 1e-6)
        ref = x + noise * (((sg + 1e-6) ** 2 - sg ** 2) ** 0.5)[:, None, None, None]
        assert torch.isfinite(xa).all() and torch.allclose(xa, ref, rtol=1e-6, atol=1e-6)
    # NCHW pack / unpack
    src = torch.randn(3, 11, h, w, device="cuda", generator=g)
    dst = torch.empty(3 * (h + 2) * (w + 2), 64, device="cuda", dtype=torch.bfloat16)
    ops.pack_nchw(src, dst)
    d4 = dst.reshape(3, h + 2, w + 2, 64)
    assert torch.equal(d4[:, 1:-1, 1:-1, :11].permute(0, 3, 1, 2), bf(src))
    assert d4[..., 11:].abs().max() == 0 and d4[:, 0].abs().max() == 0
    back = ops.unpack_nchw(eps, N, 4, h, w)
    assert torch.equal(back, e)


# ---------------------------------------------------------------------------------------------------
def test_reverse_traversal_is_bit_identical(ops):
    """`reverse` only changes the order in which tiles / rows are visited (L2-aware traversal): same bits."""
    from seva_b200 import pack

    g = torch.Generator(device="cuda").manual_seed(3)
    rnd = lambda *s: torch.randn(*s, device="cuda", generator=g)
    # linear with bias + residual, geglu, conv3x3 with row bias; several tiles per CTA and ragged tails
    A, W, b = bf(rnd(4097, 320)), bf(rnd(960, 320) / 18), rnd(960) * 0.1
    res = bf(rnd(4097, 960))
    for cl in (1, 2):
        assert torch.equal(ops.linear(A, W, b, residual=res, cluster=cl),
                           ops.linear(A, W, b, residual=res, cluster=cl, reverse=True))
    Wg, bg = bf(rnd(2560, 320) / 18), rnd(2560) * 0.1
    Wp, bp = pack.pack_geglu(Wg, bg, ops.pick_bn(2560, True))
    assert torch.equal(ops.linear(A, Wp, bp, geglu=True), ops.linear(A, Wp, bp, geglu=True, reverse=True))
    n_img, H, Cin, Cout = 3, 18, 128, 320
    ximg = torch.zeros(n_img, H + 2, H + 2, Cin, device="cuda")
    ximg[:, 1:-1, 1:-1] = rnd(n_img, H, H, Cin)
    x = bf(ximg.reshape(-1, Cin))
    W9 = pack.pack_conv3x3(bf(rnd(Cout, Cin, 3, 3) / 34))
    rb = rnd(n_img, Cout)
    assert torch.equal(ops.conv3x3_padded(x, W9, None, n_img, H, H, rowbias=rb),
                       ops.conv3x3_padded(x, W9, None, n_img, H, H, rowbias=rb, reverse=True))
    # norms
    gamma, beta = rnd(320), rnd(320)
    xn = bf(rnd(3 * 20 * 20, 320))
    dmap, dw, db = rnd(3, 18 * 18, 6), rnd(640, 6) * 0.3, rnd(640) * 0.3
    for dense in (False, True):
        kw = dict(silu=True, dense=dmap if dense else None, dense_w=dw if dense else None, dense_b=db if dense else None)
        assert torch.equal(ops.groupnorm(xn, gamma, beta, 3, 18, 18, **kw),
                           ops.groupnorm(xn, gamma, beta, 3, 18, 18, reverse=True, **kw))
    xl = bf(rnd(1001, 320))
    assert torch.equal(ops.layernorm(xl, gamma, beta), ops.layernorm(xl, gamma, beta, reverse=True))


def test_groupnorm_stats_buffer_reuse_is_deterministic(ops):
    """The ticket counters reset themselves: reusing one stats buffer across launches gives identical totals."""
    g = torch.Generator(device="cuda").manual_seed(4)
    n_img, rpi, Cc = 5, 38 * 38, 640
    x = bf(torch.randn(n_img * rpi, Cc, device="cuda", generator=g))
    sums = torch.zeros(ops.gn_stats_floats(n_img, rpi, Cc), device="cuda")
    outs = []
    for rev in (False, True, False):
        ops.groupnorm_stats(x, n_img, rpi, Cc, sums, reverse=rev)
        outs.append(sums[: n_img * Cc * 2].clone())
    assert torch.equal(outs[0], outs[1]) and torch.equal(outs[0], outs[2])
    ref = torch.stack([x.float().view(n_img, rpi, Cc).sum(1), (x.float() ** 2).view(n_img, rpi, Cc).sum(1)], -1)
    assert rel_l2(outs[0].view(n_img, Cc, 2), ref) < 1e-5


# ---------------------------------------------------------------------------------------------------
# statistics fused into the producing GEMM's epilogue (GroupNorm partials, LayerNorm row sums) and
# LayerNorm folded into the consuming GEMM
# ---------------------------------------------------------------------------------------------------
def _gn_ref_stats(x_nchw, groups, eps):
    n, c = x_nchw.shape[:2]
    xg = x_nchw.double().reshape(n, groups, -1)
    mean = xg.mean(-1)
    var = xg.var(-1, unbiased=False)
    return torch.stack([mean, 1.0 / torch.sqrt(var + eps)], -1).float()


@pytest.mark.parametrize("cluster", [1, 2])
@pytest.mark.parametrize("n_img,H,Wd,Cin,Cout,residual", [(3, 18, 18, 64, 320, True), (2, 9, 9, 128, 640, False),
                                                          (5, 2, 2, 64, 64, True), (2, 36, 20, 64, 960, False)])
def test_conv_epilogue_groupnorm_partials(ops, cluster, n_img, H, Wd, Cin, Cout, residual):
    """conv3x3 writes per-32-row-block channel partials of its own output; finalize folds them to (mean, rstd) per
    (image, group) -- compared with group statistics of the stored bf16 output, and GroupNorm through ``gstats`` with
    the two-launch path.  Images smaller than a 32-row block (2x2) and blocks straddling images are covered."""
    from seva_b200 import pack

    g = torch.Generator(device="cuda").manual_seed(H * Cout + cluster)
    x = bf(torch.randn(n_img, Cin, H, Wd, device="cuda", generator=g))
    w = torch.randn(Cout, Cin, 3, 3, device="cuda", generator=g) / math.sqrt(9 * Cin)
    b = torch.randn(Cout, device="cuda", generator=g) * 3.0  # a mean well away from zero
    xp = F.pad(x.permute(0, 2, 3, 1), (0, 0, 1, 1, 1, 1)).reshape(-1, Cin).contiguous()
    rpi = (H + 2) * (Wd + 2)
    res = None
    if residual:
        res = F.pad(bf(torch.randn(n_img, H, Wd, Cout, device="cuda", generator=g)), (0, 0, 1, 1, 1, 1)).reshape(-1, Cout).contiguous()
    part = torch.full((ops.gn_partial_floats(n_img, rpi, Cout),), float("nan"), device="cuda")
    out = ops.conv3x3_padded(xp, pack.pack_conv3x3(w), b, n_img, H, Wd, residual=res, cluster=cluster, gn_partial=part)
    plain = ops.conv3x3_padded(xp, pack.pack_conv3x3(w), b, n_img, H, Wd, residual=res, cluster=cluster)
    assert torch.equal(out, plain)  # the statistics do not perturb the output
    gst = ops.groupnorm_finalize(part, Cout, rpi, n_img, H * Wd, 1e-5)
    o = out.reshape(n_img, H + 2, Wd + 2, Cout)[:, 1:-1, 1:-1].permute(0, 3, 1, 2)
    ref = _gn_ref_stats(o.float(), 32, 1e-5)
    tol = 2e-3 if residual else 1e-5  # residual path: statistics of the fp32 values before the bf16 rounding
    assert (gst[..., 0] - ref[..., 0]).abs().max().item() < tol * (1 + ref[..., 0].abs().max().item())
    assert ((gst[..., 1] - ref[..., 1]).abs() / ref[..., 1]).max().item() < max(tol, 1e-4)
    gamma = torch.randn(Cout, device="cuda", generator=g)
    beta = torch.randn(Cout, device="cuda", generator=g)
    y_fused = ops.groupnorm(out, gamma, beta, n_img, H, Wd, silu=True, gstats=gst)
    y_two = ops.groupnorm(out, gamma, beta, n_img, H, Wd, silu=True)
    assert rel_l2(y_fused, y_two) < 2e-3
    assert torch.equal(ops.groupnorm_finalize(part, Cout, rpi, n_img, H * Wd, 1e-5), gst)  # deterministic


def test_gemm_epilogue_groupnorm_partials_two_sources(ops):
    """The decoder's virtual skip concat: two producers with DIFFERENT row spaces (a token GEMM scattered into the
    zero-bordered layout, out_map = 1, and a conv over zero-bordered rows) write column slices of one buffer; 30-channel
    groups straddle the 640 | 320 seam."""
    from seva_b200 import _lib, pack

    g = torch.Generator(device="cuda").manual_seed(77)
    n_img, H, Wd, C0, C1 = 3, 12, 10, 640, 320
    rp, rc = (H + 2) * (Wd + 2), H * Wd
    cat = torch.zeros(n_img * rp, C0 + C1, device="cuda", dtype=torch.bfloat16)
    # source 0: token GEMM -> padded rows (out_map = 1), statistics in compact row space
    A = bf(torch.randn(n_img * rc, 256, device="cuda", generator=g))
    W0 = bf(torch.randn(C0, 256, device="cuda", generator=g) / 16)
    b0 = torch.randn(C0, device="cuda", generator=g)
    part0 = torch.zeros(ops.gn_partial_floats(n_img, rc, C0), device="cuda")
    o0 = cat[:, :C0]
    gd = ops.make_gemm(A_ptr=A.data_ptr(), lda=A.stride(0), a_rows=A.shape[0], W_ptr=W0.data_ptr(), ldw=W0.stride(0),
                       M=A.shape[0], N=C0, Kc=256, out_ptr=o0.data_ptr(), ldo=o0.stride(0), bias_ptr=b0.data_ptr(),
                       out_map=1, H=H, Wd=Wd, gn_partial_ptr=part0.data_ptr(), gn_rpi=rc)
    ops.run_gemm(gd)
    # source 1: conv over zero-bordered rows
    x = bf(torch.randn(n_img, 64, H, Wd, device="cuda", generator=g))
    w = torch.randn(C1, 64, 3, 3, device="cuda", generator=g) / 24
    xp = F.pad(x.permute(0, 2, 3, 1), (0, 0, 1, 1, 1, 1)).reshape(-1, 64).contiguous()
    part1 = torch.zeros(ops.gn_partial_floats(n_img, rp, C1), device="cuda")
    ops.conv3x3_padded(xp, pack.pack_conv3x3(w), None, n_img, H, Wd, out=cat[:, C0:], gn_partial=part1)
    gst = ops.groupnorm_finalize(part0, C0, rc, n_img, rc, 1e-5, part1=part1, C1=C1, rpi1=rp)
    o = cat.reshape(n_img, H + 2, Wd + 2, C0 + C1)[:, 1:-1, 1:-1].permute(0, 3, 1, 2)
    ref = _gn_ref_stats(o.float(), 32, 1e-5)
    assert (gst[..., 0] - ref[..., 0]).abs().max().item() < 1e-5 * (1 + ref[..., 0].abs().max().item())
    assert ((gst[..., 1] - ref[..., 1]).abs() / ref[..., 1]).max().item() < 1e-4


@pytest.mark.parametrize("M,Cc,N,geglu,residual", [(1000, 320, 960, False, True), (777, 640, 1920, False, False),
                                                  (300, 1280, 3840, False, True), (20000, 320, 2560, True, True),
                                                  (515, 640, 5120, True, False), (130, 1280, 10240, True, True)])
def test_layernorm_folded_into_gemm(ops, M, Cc, N, geglu, residual):
    """producer GEMM (x = A0 W0^T + b0 [+ res]) writes per-row (sum, sumsq) partials of x; the consumer computes
    Linear(LayerNorm(x)) [GEGLU] with gamma folded into W, beta into the bias and mean / rstd applied in the epilogue.
    Reference: fp32 LayerNorm + Linear on the stored bf16 x."""
    from seva_b200 import pack

    g = torch.Generator(device="cuda").manual_seed(M + N)
    A0 = bf(torch.randn(M, 256, device="cuda", generator=g))
    W0 = bf(torch.randn(Cc, 256, device="cuda", generator=g) / 16)
    b0 = torch.randn(Cc, device="cuda", generator=g) * 2 + 1.0
    res = bf(torch.randn(M, Cc, device="cuda", generator=g)) if residual else None
    rows = torch.full((ops.ln_slots(Cc), M, 2), float("nan"), device="cuda")
    x = ops.linear(A0, W0, b0, residual=res, ln_rowstats=rows)
    assert torch.equal(x, ops.linear(A0, W0, b0, residual=res))
    s_ref = x.float().sum(-1)
    assert (rows[..., 0].sum(0) - s_ref).abs().max().item() < 2e-3 * x.float().abs().sum(-1).max().item()
    gamma = 1 + 0.3 * torch.randn(Cc, device="cuda", generator=g)
    beta = 0.3 * torch.randn(Cc, device="cuda", generator=g)
    W = torch.randn(N, Cc, device="cuda", generator=g) / math.sqrt(Cc)
    b = torch.randn(N, device="cuda", generator=g) * 0.1
    Wg = (W * gamma[None]).to(torch.bfloat16)
    colsum, bias2 = Wg.float().sum(1), W @ beta + b
    if geglu:
        bn = ops.pick_bn(N, True)
        Wp, bp = pack.pack_geglu(Wg, bias2, bn)
        _, cp = pack.pack_geglu(Wg, colsum, bn)
        out = ops.linear(x, Wp, bp, geglu=True, bn=bn, ln_in=rows, ln_colsum=cp)
    else:
        out = ops.linear(x, Wg, bias2.contiguous(), ln_in=rows, ln_colsum=colsum.contiguous())
    y = F.layer_norm(x.float(), (Cc,), gamma, beta, 1e-5) @ W.t() + b
    if geglu:
        a_, gate = y.chunk(2, dim=-1)
        y = a_ * F.gelu(gate)
    assert rel_l2(out, y) < 6e-3, rel_l2(out, y)


@pytest.mark.parametrize("fused", [False, True])
def test_groupnorm_large_mean(ops, fused):
    """groups with |mean| >> std (activations of a trained checkpoint can sit at mean / std ~ 25-100).
    The sums are fp32; the E[x^2] - mean^2 step is done in fp64 (apply prologue / finalize)."""
    from seva_b200 import pack

    g = torch.Generator(device="cuda").manual_seed(42)
    n_img, H, Wd, Cc = 2, 36, 36, 320
    if fused:  # statistics from a conv epilogue whose bias puts the output at mean ~ 100, std ~ 1
        x = bf(torch.randn(n_img, 64, H, Wd, device="cuda", generator=g))
        w = torch.randn(Cc, 64, 3, 3, device="cuda", generator=g) / 24
        b = torch.full((Cc,), 100.0, device="cuda") + torch.randn(Cc, device="cuda", generator=g)
        xp = F.pad(x.permute(0, 2, 3, 1), (0, 0, 1, 1, 1, 1)).reshape(-1, 64).contiguous()
        rpi = (H + 2) * (Wd + 2)
        part = torch.zeros(ops.gn_partial_floats(n_img, rpi, Cc), device="cuda")
        xin = ops.conv3x3_padded(xp, pack.pack_conv3x3(w), b, n_img, H, Wd, gn_partial=part)
        gst = ops.groupnorm_finalize(part, Cc, rpi, n_img, H * Wd, 1e-5)
    else:
        xc = bf(torch.randn(n_img, H, Wd, Cc, device="cuda", generator=g) * 4.0 + 100.0)
        xin = F.pad(xc, (0, 0, 1, 1, 1, 1)).reshape(-1, Cc).contiguous()
        gst = None
    gamma = torch.randn(Cc, device="cuda", generator=g)
    beta = torch.randn(Cc, device="cuda", generator=g)
    out = ops.groupnorm(xin, gamma, beta, n_img, H, Wd, silu=False, gstats=gst)
    xs = xin.reshape(n_img, H + 2, Wd + 2, Cc)[:, 1:-1, 1:-1].permute(0, 3, 1, 2).float()
    ref = F.group_norm(xs.double(), 32, gamma.double(), beta.double(), 1e-5).float()
    got = out.reshape(n_img, H + 2, Wd + 2, Cc)[:, 1:-1, 1:-1].permute(0, 3, 1, 2)
    assert rel_l2(got, ref) < 4e-3, rel_l2(got, ref)


@pytest.mark.parametrize("B,heads,Lq,blocks", [(2, 5, 333, (300, 77, 411)), (1, 10, 1000, (128, 128)), (2, 2, 70, (1, 500, 64, 3))])
def test_attention_partial_and_merge(ops, B, heads, Lq, blocks):
    """Key blocks attended one launch at a time (the frame-shard ring): fp32 partial results + log-sum-exp, merged,
    equal one launch over all keys."""
    g = torch.Generator(device="cuda").manual_seed(Lq)
    Cc = heads * 64
    Lk = sum(blocks)
    q = bf(torch.randn(B, Lq, Cc, device="cuda", generator=g))
    k = bf(torch.randn(B, Lk, Cc, device="cuda", generator=g))
    v = bf(torch.randn(B, Lk, Cc, device="cuda", generator=g))
    full = ops.attention(q, k, v, heads)
    parts = torch.empty(len(blocks), B, Lq, Cc, device="cuda")
    lse = torch.empty(len(blocks), B, heads, Lq, device="cuda")
    o = 0
    for i, n in enumerate(blocks):
        ops.attention_partial(q, k[:, o:o + n], v[:, o:o + n], heads, out=parts[i], lse=lse[i])
        o += n
    merged = ops.attention_merge(parts, lse)
    sp = lambda t: t.reshape(B, -1, heads, 64).permute(0, 2, 1, 3).float()
    ref = F.scaled_dot_product_attention(sp(q), sp(k), sp(v)).permute(0, 2, 1, 3).reshape(B, Lq, Cc)
    assert rel_l2(merged, ref) < 5e-3, rel_l2(merged, ref)
    assert rel_l2(merged, full) < 3e-3, rel_l2(merged, full)
