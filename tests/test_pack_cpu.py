"""Host logic: the packed weight layouts of seva_b200/pack.py against what the kernels do with them, restated in plain
torch on the CPU (no CUDA): the nine row-shifted taps of the implicit-GEMM convolution over a zero-bordered flat image, the
four 2 x 2 phase convolutions that replace nearest-2x upsampling + 3 x 3 convolution (SevaGemm ntaps = 4, out_map = 2..5,
include/seva_b200.h), and the per-N-tile [x | gate] interleave the GEGLU epilogue expects.  The weights are compared as the
kernels see them (bf16-rounded), so the tolerances only cover fp32 summation order -- except for the phase convolutions,
whose tap SUMS are rounded to bf16 once (pack_upconv_phases) where the reference rounds each tap.
"""
import os
import sys

import torch
import torch.nn.functional as F

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "stable-virtual-camera_b200")]

from seva_b200 import pack  # noqa: E402


def _rel(a, b):
    return ((a.double() - b.double()).norm() / b.double().norm()).item()


def _padded_flat(x):
    """[n, C, H, W] -> zero-bordered channels-last flat rows [n * (H + 2) * (W + 2), C]"""
    n, c, h, w = x.shape
    return F.pad(x, (1, 1, 1, 1)).permute(0, 2, 3, 1).reshape(n * (h + 2) * (w + 2), c)


def _shifted(rows, off):
    """rows[m + off] with rows outside the tensor read as zero (TMA zero fill)"""
    out = torch.zeros_like(rows)
    M = rows.shape[0]
    lo, hi = max(0, -off), min(M, M - off)
    out[lo:hi] = rows[lo + off:hi + off]
    return out


def test_conv3x3_is_nine_row_shifted_gemm_taps():
    g = torch.Generator().manual_seed(0)
    n, cin, cout, h, w = 2, 24, 40, 6, 9
    x = torch.randn(n, cin, h, w, generator=g)
    wt = torch.randn(cout, cin, 3, 3, generator=g) / (9 * cin) ** 0.5
    W9 = pack.pack_conv3x3(wt)
    assert W9.shape == (cout, 9 * cin) and W9.dtype == torch.bfloat16
    rows = _padded_flat(x)
    wp = w + 2
    acc = torch.zeros(rows.shape[0], cout)
    for tap in range(9):   # tap (dy, dx) of output row m is input row m + dy * Wp + dx
        off = (tap // 3 - 1) * wp + (tap % 3 - 1)
        acc += _shifted(rows, off) @ W9[:, tap * cin:(tap + 1) * cin].float().t()
    got = acc.reshape(n, h + 2, wp, cout)[:, 1:-1, 1:-1].permute(0, 3, 1, 2)   # the kernel zeroes the border rows itself
    ref = F.conv2d(x, wt.bfloat16().float(), padding=1)
    assert _rel(got, ref) < 1e-6
    # padded input channels (conv_in: 11 -> 64) are zero columns
    W9p = pack.pack_conv3x3(wt, cin_pad=32)
    assert W9p.shape == (cout, 9 * 32)
    v = W9p.view(cout, 9, 32)
    assert torch.equal(v[:, :, :cin].reshape(cout, -1), W9) and not v[:, :, cin:].any()


def test_upsample_conv_equals_four_phase_convs():
    g = torch.Generator().manual_seed(1)
    n, cin, cout, h, w = 2, 16, 24, 5, 7
    x = torch.randn(n, cin, h, w, generator=g)
    wt = torch.randn(cout, cin, 3, 3, generator=g) / (9 * cin) ** 0.5
    phases = pack.pack_upconv_phases(wt)
    assert len(phases) == 4 and all(p.shape == (cout, 4 * cin) and p.dtype == torch.bfloat16 for p in phases)
    rows = _padded_flat(x)                      # the LOW-resolution zero-bordered image is the A operand
    wp = w + 2
    out = torch.zeros(n, 2 * h + 2, 2 * w + 2, cout)
    for ph in range(4):
        py, px = ph >> 1, ph & 1
        acc = torch.zeros(rows.shape[0], cout)
        for tap in range(4):                    # row m + (tap / 2 + py - 1) * Wp + (tap % 2 + px - 1)
            off = ((tap >> 1) + py - 1) * wp + ((tap & 1) + px - 1)
            acc += _shifted(rows, off) @ phases[ph][:, tap * cin:(tap + 1) * cin].float().t()
        acc = acc.reshape(n, h + 2, wp, cout)
        for Y in range(1, h + 1):               # interior pixel (Y, X) -> pixel (2Y - 1 + py, 2X - 1 + px) of the 2x image
            for X in range(1, w + 1):
                out[:, 2 * Y - 1 + py, 2 * X - 1 + px] = acc[:, Y, X]
    got = out[:, 1:-1, 1:-1].permute(0, 3, 1, 2)
    ref = F.conv2d(F.interpolate(x, scale_factor=2, mode="nearest"), wt, padding=1)
    assert not out[:, 0].any() and not out[:, -1].any() and not out[:, :, 0].any() and not out[:, :, -1].any()
    assert _rel(got, ref) < 4e-3                # bf16 rounding of the summed taps
    # with weights that survive the rounding the identity is exact up to summation order
    wq = torch.randint(-4, 5, (cout, cin, 3, 3), generator=g).float() / 8
    ph_q = pack.pack_upconv_phases(wq)
    acc = torch.zeros(rows.shape[0], cout)
    for tap in range(4):
        acc += _shifted(rows, ((tap >> 1) + 1 - 1) * wp + ((tap & 1) + 0 - 1)) @ ph_q[2][:, tap * cin:(tap + 1) * cin].float().t()
    ref_q = F.conv2d(F.interpolate(x, scale_factor=2, mode="nearest"), wq, padding=1)
    got_q = acc.reshape(n, h + 2, wp, cout)[:, 1:-1, 1:-1].permute(0, 3, 1, 2)     # phase (1, 0): rows 2Y, columns 2X - 1
    assert _rel(got_q, ref_q[:, :, 1::2, 0::2]) < 1e-6


def test_geglu_rows_are_interleaved_per_n_tile():
    g = torch.Generator().manual_seed(2)
    inner, K, M = 192, 48, 10
    for bn in (64, 128, 192):
        w = torch.randn(2 * inner, K, generator=g) / K ** 0.5
        b = torch.randn(2 * inner, generator=g) * 0.1
        a = torch.randn(M, K, generator=g)
        Wp, bp = pack.pack_geglu(w, b, bn)
        assert Wp.shape == w.shape and Wp.dtype == torch.bfloat16 and bp.dtype == torch.float32
        y = a @ Wp.float().t() + bp              # what the accumulator tiles hold, tile t = columns [t * bn, (t + 1) * bn)
        out = torch.empty(M, inner)
        half = bn // 2
        for t in range(2 * inner // bn):         # epilogue: out[:, t * bn / 2 + j] = x_j * gelu(gate_j)
            tile = y[:, t * bn:(t + 1) * bn]
            out[:, t * half:(t + 1) * half] = tile[:, :half] * F.gelu(tile[:, half:])
        x_, gate = (a @ w.bfloat16().float().t() + b).chunk(2, dim=-1)   # GEGLU.forward (transformer.py:14)
        assert _rel(out, x_ * F.gelu(gate)) < 1e-6
        perm = pack.geglu_perm(inner, bn)
        assert sorted(perm.tolist()) == list(range(2 * inner))


def test_linear_and_1x1_weights_are_k_major_bf16():
    w = torch.randn(12, 20)
    assert torch.equal(pack.pack_linear(w), w.bfloat16()) and pack.pack_linear(w).is_contiguous()
    w4 = torch.randn(12, 20, 1, 1)
    assert torch.equal(pack.pack_conv1x1(w4), w4.reshape(12, 20).bfloat16())


# ---- engine-side packing: zero-padded narrow heads and the LayerNorm fold (host logic of engine.pack_weights) ----------
class _Attn(torch.nn.Module):
    def __init__(self, dim, inner):
        super().__init__()
        self.to_q, self.to_k, self.to_v = (torch.nn.Linear(dim, inner, bias=False) for _ in range(3))
        self.to_out = torch.nn.ModuleList([torch.nn.Linear(inner, dim)])


def test_narrow_heads_zero_padded_to_the_kernel_width():
    """num_head_channels < 64: q / k / v rows and to_out columns are padded per head to 64; with the softmax scale of the
    TRUE width the padded attention is the original one (the zero dimensions add nothing to q.k and produce zero outputs
    that meet zero columns of to_out)."""
    from seva_b200.engine import _pack_attn_self, _pad_head_cols, _pad_head_rows

    torch.manual_seed(3)
    for heads, dh in ((3, 16), (2, 32), (5, 8), (2, 64)):
        dim, L = 40, 11
        at = _Attn(dim, heads * dh)
        p = _pack_attn_self(at, heads, dh)
        assert p["wqkv"].shape == (3 * heads * 64, dim) and p["wo"].shape == (dim, heads * 64)
        x = torch.randn(L, dim)
        q, k, v = (x @ p["wqkv"].float().t()).view(L, 3, heads, 64).permute(1, 2, 0, 3)     # [heads, L, 64] each
        o = torch.softmax(q @ k.transpose(-1, -2) / dh ** 0.5, -1) @ v
        got = o.permute(1, 0, 2).reshape(L, heads * 64) @ p["wo"].float().t() + p["bo"]
        r = lambda w: w.detach().bfloat16().float()
        q0, k0, v0 = ((x @ r(m.weight).t()).view(L, heads, dh).transpose(0, 1) for m in (at.to_q, at.to_k, at.to_v))
        o0 = F.scaled_dot_product_attention(q0[None], k0[None], v0[None])[0]
        ref = o0.transpose(0, 1).reshape(L, heads * dh) @ r(at.to_out[0].weight).t() + at.to_out[0].bias
        assert _rel(got, ref) < 1e-6
        w = torch.randn(heads * dh, 7)
        back = _pad_head_rows(w, heads, dh).view(heads, 64, 7)
        assert torch.equal(back[:, :dh].reshape(heads * dh, 7), w) and not back[:, dh:].any()
        assert torch.equal(_pad_head_cols(w.t(), heads, dh), _pad_head_rows(w, heads, dh).t())


def test_layernorm_fold_identity():
    """Linear(LayerNorm(x)) = rstd * (x W'^T - mean * colsum) + bias' with W' = W * gamma (bf16), colsum = sum_k W'[n, k],
    bias' = W beta + b -- the form the GEMM epilogue applies (SevaGemm.ln_in_stats / ln_colsum)."""
    from seva_b200.engine import _fold_ln_linear

    torch.manual_seed(4)
    C, N, M = 48, 20, 9
    ln = torch.nn.LayerNorm(C)
    with torch.no_grad():
        ln.weight.copy_(1 + 0.2 * torch.randn(C))
        ln.bias.copy_(0.1 * torch.randn(C))
    w, b, x = torch.randn(N, C) / C ** 0.5, torch.randn(N) * 0.1, torch.randn(M, C) * 2 + 0.5
    wg, colsum, bias2 = _fold_ln_linear(ln, w, b)
    assert wg.dtype == torch.bfloat16
    mean, var = x.mean(1, keepdim=True), x.var(1, unbiased=False, keepdim=True)
    rstd = torch.rsqrt(var + ln.eps)
    got = rstd * (x @ wg.float().t() - mean * colsum[None]) + bias2
    gamma_w = wg.float() / ln.weight.detach()[None]          # the weight the fold effectively applies after rounding
    ref = F.linear(F.layer_norm(x, (C,), ln.weight, ln.bias, ln.eps), gamma_w, b) - F.linear(ln.bias.detach()[None], gamma_w) \
        + F.linear(ln.bias.detach()[None], w)
    assert _rel(got, ref) < 1e-5
    assert _rel(got, F.linear(F.layer_norm(x, (C,), ln.weight, ln.bias, ln.eps), w, b)) < 5e-3   # bf16 rounding of W * gamma
