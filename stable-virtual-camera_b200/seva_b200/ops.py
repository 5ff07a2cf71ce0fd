"""Tensor-level wrappers over the C ABI (``include/seva_b200.h``).

torch is used only for device memory and the current stream; every function launches one of our
sm_100a kernels.  CPU tensors are rejected: there is no fallback path.
"""
from __future__ import annotations

import ctypes as C
import math

import torch

from . import _lib
from ._lib import SevaGemm, SevaGNApply, check


def _stream() -> int:
    return torch.cuda.current_stream().cuda_stream


def _p(t: torch.Tensor | None) -> int | None:
    return None if t is None else t.data_ptr()


def _req(t: torch.Tensor, dtype, name: str) -> torch.Tensor:
    if not t.is_cuda:
        raise _lib.SevaB200Error(f"{name} must be a CUDA tensor (seva_b200 has no CPU fallback)")
    if t.dtype != dtype:
        raise _lib.SevaB200Error(f"{name} must be {dtype}, got {t.dtype}")
    return t


def check_device() -> None:
    check(_lib.load().seva_check_device(), "seva_check_device")


def pick_bn(N: int, geglu: bool = False) -> int:
    return _lib.load().seva_gemm_pick_bn(N, int(geglu))


# ------------------------------------------------------------------------------------------------
# GEMM / conv
# ------------------------------------------------------------------------------------------------
def make_gemm(*, A_ptr, lda, a_rows, W_ptr, ldw, M, N, Kc, out_ptr, ldo, ntaps=1, tap_wp=0, bias_ptr=None,
              rowbias_ptr=None, rowbias_rpi=0, residual_ptr=None, ldr=0, out_fp32=False, geglu=False,
              out_map=0, H=0, Wd=0, mask_border=False, Hp=0, Wp=0, bn=0, cluster=0, reverse=False,
              gn_partial_ptr=None, gn_rpi=0, ln_rowstats_ptr=None, ln_in_ptr=None, ln_in_slots=0, ln_eps=1e-5,
              ln_colsum_ptr=None) -> SevaGemm:
    g = SevaGemm()
    g.A, g.lda, g.a_rows = A_ptr, lda, a_rows
    g.W, g.ldw = W_ptr, ldw
    g.M, g.N, g.Kc, g.ntaps, g.tap_wp = M, N, Kc, ntaps, tap_wp
    g.bias, g.rowbias, g.rowbias_rpi = bias_ptr, rowbias_ptr, rowbias_rpi
    g.residual, g.ldr = residual_ptr, ldr
    g.out, g.ldo = out_ptr, ldo
    g.out_fp32, g.geglu, g.out_map, g.H, g.Wd = int(out_fp32), int(geglu), out_map, H, Wd
    g.mask_border, g.Hp, g.Wp, g.bn, g.cluster = int(mask_border), Hp, Wp, bn, cluster
    g.reverse = int(reverse)
    g.gn_partial, g.gn_rpi, g.gn_slots = gn_partial_ptr, gn_rpi, ((gn_rpi + 31) // 32 + 1) if gn_partial_ptr else 0
    g.ln_rowstats = ln_rowstats_ptr
    g.ln_in_stats, g.ln_in_slots, g.ln_eps, g.ln_colsum = ln_in_ptr, ln_in_slots, ln_eps, ln_colsum_ptr
    return g


def ln_slots(N: int, bn: int = 0) -> int:
    """Row-statistic slots a GEMM with N output columns writes per row (SevaGemm.ln_rowstats)."""
    return _lib.load().seva_gemm_ln_slots(N, bn)


def gn_partial_floats(n_img: int, rpi: int, Cc: int) -> int:
    n = _lib.load().seva_groupnorm_partial_floats(n_img, rpi, Cc)
    if n < 0:
        raise _lib.SevaB200Error(f"seva_groupnorm_partial_floats({n_img}, {rpi}, {Cc}) failed")
    return int(n)


def groupnorm_finalize(part0: torch.Tensor, C0: int, rpi0: int, n_img: int, pixels: int, eps: float, *, groups: int = 32,
                       part1: torch.Tensor | None = None, C1: int = 0, rpi1: int = 0,
                       out: torch.Tensor | None = None) -> torch.Tensor:
    """Fold GEMM-epilogue partials into gstats [n_img, groups, 2] = (mean, rstd)."""
    if out is None:
        out = torch.empty((n_img, groups, 2), device=part0.device, dtype=torch.float32)
    check(_lib.load().seva_groupnorm_finalize(part0.data_ptr(), C0, rpi0, _p(part1), C1, rpi1, n_img, groups, pixels,
                                              eps, out.data_ptr(), _stream()), "seva_groupnorm_finalize")
    return out


def run_gemm(g: SevaGemm) -> None:
    check(_lib.load().seva_gemm_bf16(C.byref(g), _stream()), "seva_gemm_bf16")


def linear(A: torch.Tensor, W: torch.Tensor, bias: torch.Tensor | None = None, *, out: torch.Tensor | None = None,
           residual: torch.Tensor | None = None, rowbias: torch.Tensor | None = None, rowbias_rpi: int = 0,
           geglu: bool = False, out_fp32: bool = False, bn: int = 0, cluster: int = 0,
           reverse: bool = False, gn_partial: torch.Tensor | None = None, gn_rpi: int = 0,
           ln_rowstats: torch.Tensor | None = None, ln_in: torch.Tensor | None = None, ln_eps: float = 1e-5,
           ln_colsum: torch.Tensor | None = None) -> torch.Tensor:
    """out = A @ W.T (+bias) (+rowbias[row // rpi]) (+residual); A [M,K] bf16, W [N,K] bf16 (row strides free).
    ``gn_partial`` / ``ln_rowstats``: statistics of the output for a following GroupNorm / LayerNorm;
    ``ln_in`` ([slots, M, 2] from the producer of A) + ``ln_colsum``: LayerNorm(A) folded into this GEMM."""
    _req(A, torch.bfloat16, "A"), _req(W, torch.bfloat16, "W")
    M, K = A.shape
    N = W.shape[0]
    n_out = N // 2 if geglu else N
    if out is None:
        out = torch.empty((M, n_out), device=A.device, dtype=torch.float32 if out_fp32 else torch.bfloat16)
    g = make_gemm(A_ptr=A.data_ptr(), lda=A.stride(0), a_rows=M, W_ptr=W.data_ptr(), ldw=W.stride(0), M=M, N=N,
                  Kc=K, out_ptr=out.data_ptr(), ldo=out.stride(0), bias_ptr=_p(bias), rowbias_ptr=_p(rowbias),
                  rowbias_rpi=rowbias_rpi, residual_ptr=_p(residual),
                  ldr=residual.stride(0) if residual is not None else 0, out_fp32=out_fp32, geglu=geglu, bn=bn,
                  cluster=cluster, reverse=reverse, gn_partial_ptr=_p(gn_partial), gn_rpi=gn_rpi,
                  ln_rowstats_ptr=_p(ln_rowstats), ln_in_ptr=_p(ln_in), ln_in_slots=0 if ln_in is None else ln_in.shape[0],
                  ln_eps=ln_eps, ln_colsum_ptr=_p(ln_colsum))
    run_gemm(g)
    return out


def conv3x3_padded(x: torch.Tensor, W9: torch.Tensor, bias: torch.Tensor | None, N_img: int, H: int, Wd: int, *,
                   out: torch.Tensor | None = None, residual: torch.Tensor | None = None,
                   rowbias: torch.Tensor | None = None, out_fp32: bool = False, bn: int = 0,
                   cluster: int = 0, reverse: bool = False, gn_partial: torch.Tensor | None = None) -> torch.Tensor:
    """3x3/s1/p1 conv on a zero-bordered flat image x [N_img*(H+2)*(Wd+2), Cin]; W9 [Cout, 9*Cin] tap-major."""
    _req(x, torch.bfloat16, "x"), _req(W9, torch.bfloat16, "W9")
    rows, Cin = x.shape
    Hp, Wp = H + 2, Wd + 2
    assert rows == N_img * Hp * Wp and W9.shape[1] == 9 * Cin
    Cout = W9.shape[0]
    if out is None:
        out = torch.empty((rows, Cout), device=x.device, dtype=torch.float32 if out_fp32 else torch.bfloat16)
    g = make_gemm(A_ptr=x.data_ptr(), lda=x.stride(0), a_rows=rows, W_ptr=W9.data_ptr(), ldw=W9.stride(0), M=rows,
                  N=Cout, Kc=Cin, ntaps=9, tap_wp=Wp, out_ptr=out.data_ptr(), ldo=out.stride(0), bias_ptr=_p(bias),
                  rowbias_ptr=_p(rowbias), rowbias_rpi=Hp * Wp, residual_ptr=_p(residual),
                  ldr=residual.stride(0) if residual is not None else 0, out_fp32=out_fp32, mask_border=True,
                  Hp=Hp, Wp=Wp, bn=bn, cluster=cluster, reverse=reverse, gn_partial_ptr=_p(gn_partial), gn_rpi=Hp * Wp)
    run_gemm(g)
    return out


# ------------------------------------------------------------------------------------------------
# attention
# ------------------------------------------------------------------------------------------------
def attention_qkv(qkv: torch.Tensor, B: int, L: int, heads: int, out: torch.Tensor | None = None) -> torch.Tensor:
    """Self-attention on a fused projection qkv [B*L, 3*heads*64] -> [B*L, heads*64]."""
    _req(qkv, torch.bfloat16, "qkv")
    Cc = heads * 64
    assert qkv.shape == (B * L, 3 * Cc)
    if out is None:
        out = torch.empty((B * L, Cc), device=qkv.device, dtype=torch.bfloat16)
    ld = qkv.stride(0)
    base = qkv.data_ptr()
    check(_lib.load().seva_attn_fwd(base, base + 2 * Cc, base + 4 * Cc, out.data_ptr(), B, heads, L, L, ld, ld, ld,
                                    out.stride(0), L * ld, L * ld, L * ld, L * out.stride(0),
                                    1.0 / math.sqrt(64.0), _stream()), "seva_attn_fwd")
    return out


def attention(q: torch.Tensor, k: torch.Tensor, v: torch.Tensor, heads: int,
              out: torch.Tensor | None = None) -> torch.Tensor:
    """q [B, Lq, heads*64], k/v [B, Lk, heads*64] (last dim contiguous) -> [B, Lq, heads*64]."""
    for t, n in ((q, "q"), (k, "k"), (v, "v")):
        _req(t, torch.bfloat16, n)
        assert t.stride(2) == 1
    B, Lq, Cc = q.shape
    Lk = k.shape[1]
    if out is None:
        out = torch.empty((B, Lq, Cc), device=q.device, dtype=torch.bfloat16)
    check(_lib.load().seva_attn_fwd(q.data_ptr(), k.data_ptr(), v.data_ptr(), out.data_ptr(), B, heads, Lq, Lk,
                                    q.stride(1), k.stride(1), v.stride(1), out.stride(1), q.stride(0), k.stride(0),
                                    v.stride(0), out.stride(0), 1.0 / math.sqrt(64.0), _stream()), "seva_attn_fwd")
    return out


def attention_partial(q: torch.Tensor, k: torch.Tensor, v: torch.Tensor, heads: int, out: torch.Tensor | None = None,
                      lse: torch.Tensor | None = None):
    """Attention of q [B, Lq, heads*64] over ONE key block k / v [B, Lk, heads*64]: (fp32 result [B, Lq, heads*64],
    base-2 log-sum-exp [B, heads, Lq]) for ``attention_merge``."""
    for t, n in ((q, "q"), (k, "k"), (v, "v")):
        _req(t, torch.bfloat16, n)
        assert t.stride(2) == 1
    B, Lq, Cc = q.shape
    Lk = k.shape[1]
    if out is None:
        out = torch.empty((B, Lq, Cc), device=q.device, dtype=torch.float32)
    if lse is None:
        lse = torch.empty((B, heads, Lq), device=q.device, dtype=torch.float32)
    check(_lib.load().seva_attn_partial(q.data_ptr(), k.data_ptr(), v.data_ptr(), out.data_ptr(), lse.data_ptr(), B, heads,
                                        Lq, Lk, q.stride(1), k.stride(1), v.stride(1), q.stride(0), k.stride(0),
                                        v.stride(0), 1.0 / math.sqrt(64.0), _stream()), "seva_attn_partial")
    return out, lse


def attention_merge(partials: torch.Tensor, lse: torch.Tensor, out: torch.Tensor | None = None) -> torch.Tensor:
    """partials [P, B, Lq, C] fp32, lse [P, B, heads, Lq] fp32 -> bf16 [B, Lq, C]."""
    _req(partials, torch.float32, "partials"), _req(lse, torch.float32, "lse")
    assert partials.is_contiguous() and lse.is_contiguous()
    P, B, Lq, Cc = partials.shape
    heads = lse.shape[2]
    if out is None:
        out = torch.empty((B, Lq, Cc), device=partials.device, dtype=torch.bfloat16)
    check(_lib.load().seva_attn_merge(partials.data_ptr(), lse.data_ptr(), out.data_ptr(), P, B, heads, Lq, out.stride(1),
                                      out.stride(0), _stream()), "seva_attn_merge")
    return out


def timemix_attention(qkv: torch.Tensor, B: int, T: int, S: int, heads: int,
                      out: torch.Tensor | None = None) -> torch.Tensor:
    _req(qkv, torch.bfloat16, "qkv")
    Cc = heads * 64
    assert qkv.shape == (B * T * S, 3 * Cc)
    if out is None:
        out = torch.empty((B * T * S, Cc), device=qkv.device, dtype=torch.bfloat16)
    check(_lib.load().seva_timemix_attn(qkv.data_ptr(), out.data_ptr(), B, T, S, heads, qkv.stride(0),
                                        out.stride(0), 1.0 / math.sqrt(64.0), _stream()), "seva_timemix_attn")
    return out


def timemix_attention_kv(q: torch.Tensor, kv: torch.Tensor, B: int, Tq: int, Tk: int, S: int, heads: int,
                         out: torch.Tensor | None = None) -> torch.Tensor:
    """Frame-axis attention with Tq local query frames against the Tk frames of ``kv`` = [k | v] (frame
    sequence-parallel ranks)."""
    _req(q, torch.bfloat16, "q")
    _req(kv, torch.bfloat16, "kv")
    Cc = heads * 64
    assert q.shape == (B * Tq * S, Cc) and kv.shape == (B * Tk * S, 2 * Cc)
    if out is None:
        out = torch.empty((B * Tq * S, Cc), device=q.device, dtype=torch.bfloat16)
    check(_lib.load().seva_timemix_attn_kv(q.data_ptr(), kv.data_ptr(), kv.data_ptr() + 2 * Cc, out.data_ptr(), B, Tq,
                                           Tk, S, heads, q.stride(0), kv.stride(0), out.stride(0),
                                           1.0 / math.sqrt(64.0), _stream()), "seva_timemix_attn_kv")
    return out


# ------------------------------------------------------------------------------------------------
# norms
# ------------------------------------------------------------------------------------------------
def gn_stats_floats(n_img: int, rpi: int, Cc: int) -> int:
    n = _lib.load().seva_groupnorm_stats_floats(n_img, rpi, Cc)
    if n < 0:
        raise _lib.SevaB200Error(f"seva_groupnorm_stats_floats({n_img}, {rpi}, {Cc}) failed")
    return int(n)


def groupnorm_stats(x: torch.Tensor, n_img: int, rpi: int, Cc: int, sums: torch.Tensor | None = None,
                    reverse: bool = False) -> torch.Tensor:
    """``sums``: a ZERO-initialised fp32 buffer of ``gn_stats_floats(n_img, rpi, Cc)`` elements (reusable across
    launches of the same shape); the [n_img, Cc, 2] totals are its first elements."""
    _req(x, torch.bfloat16, "x")
    if sums is None:
        sums = torch.zeros(gn_stats_floats(n_img, rpi, Cc), device=x.device, dtype=torch.float32)
    check(_lib.load().seva_groupnorm_stats(x.data_ptr(), x.stride(0), n_img, rpi, Cc, sums.data_ptr(), int(reverse),
                                           _stream()),
          "seva_groupnorm_stats")
    return sums


def make_gn_apply(*, x_ptr, ldx, y_ptr, ldy, sums_ptr, gamma_ptr, beta_ptr, n_img, H, Wd, Cc, groups, eps,
                  in_padded, out_padded, silu, dense_ptr=None, dense_w_ptr=None, dense_b_ptr=None,
                  reverse=False, gstats_ptr=None) -> SevaGNApply:
    p = SevaGNApply()
    p.x, p.ldx, p.y, p.ldy = x_ptr, ldx, y_ptr, ldy
    p.sums, p.gamma, p.beta = sums_ptr, gamma_ptr, beta_ptr
    p.Nimg, p.H, p.Wd, p.C, p.groups = n_img, H, Wd, Cc, groups
    p.eps, p.in_padded, p.out_padded, p.silu = eps, int(in_padded), int(out_padded), int(silu)
    p.dense, p.dense_w, p.dense_b = dense_ptr, dense_w_ptr, dense_b_ptr
    p.reverse = int(reverse)
    p.gstats = gstats_ptr
    return p


def run_gn_apply(p: SevaGNApply) -> None:
    check(_lib.load().seva_groupnorm_apply(C.byref(p), _stream()), "seva_groupnorm_apply")


def groupnorm(x: torch.Tensor, gamma: torch.Tensor, beta: torch.Tensor, n_img: int, H: int, Wd: int, *, groups=32,
              eps=1e-5, in_padded=True, out_padded=True, silu=False, dense=None, dense_w=None, dense_b=None,
              out: torch.Tensor | None = None, reverse: bool = False,
              gstats: torch.Tensor | None = None) -> torch.Tensor:
    """GroupNorm (+SiLU, +dense modulation) of a channels-last image batch (two launches: stats, apply; with
    ``gstats`` from ``groupnorm_finalize`` only the apply launch)."""
    _req(x, torch.bfloat16, "x")
    Cc = x.shape[1]
    rpi_in = (H + 2) * (Wd + 2) if in_padded else H * Wd
    rpi_out = (H + 2) * (Wd + 2) if out_padded else H * Wd
    assert x.shape[0] == n_img * rpi_in
    sums = None if gstats is not None else groupnorm_stats(x, n_img, rpi_in, Cc, reverse=reverse)
    if out is None:
        out = torch.empty((n_img * rpi_out, Cc), device=x.device, dtype=torch.bfloat16)
    p = make_gn_apply(x_ptr=x.data_ptr(), ldx=x.stride(0), y_ptr=out.data_ptr(), ldy=out.stride(0),
                      sums_ptr=_p(sums), gamma_ptr=gamma.data_ptr(), beta_ptr=beta.data_ptr(), n_img=n_img,
                      H=H, Wd=Wd, Cc=Cc, groups=groups, eps=eps, in_padded=in_padded, out_padded=out_padded,
                      silu=silu, dense_ptr=_p(dense), dense_w_ptr=_p(dense_w), dense_b_ptr=_p(dense_b),
                      reverse=not reverse, gstats_ptr=_p(gstats))
    run_gn_apply(p)
    return out


def layernorm(x: torch.Tensor, gamma: torch.Tensor, beta: torch.Tensor, eps: float = 1e-5,
              out: torch.Tensor | None = None, reverse: bool = False) -> torch.Tensor:
    _req(x, torch.bfloat16, "x"), _req(gamma, torch.float32, "gamma"), _req(beta, torch.float32, "beta")
    M, Cc = x.shape
    if out is None:
        out = torch.empty((M, Cc), device=x.device, dtype=torch.bfloat16)
    check(_lib.load().seva_layernorm(x.data_ptr(), x.stride(0), out.data_ptr(), out.stride(0), gamma.data_ptr(),
                                     beta.data_ptr(), M, Cc, eps, int(reverse), _stream()), "seva_layernorm")
    return out


# ------------------------------------------------------------------------------------------------
# geometry
# ------------------------------------------------------------------------------------------------
def plucker(w2c_src: torch.Tensor, w2c: torch.Tensor, Kpix: torch.Tensor, h: int, w: int) -> torch.Tensor:
    for t, n in ((w2c_src, "w2c_src"), (w2c, "w2c"), (Kpix, "Kpix")):
        _req(t, torch.float32, n)
        assert t.is_contiguous()
    T = w2c.shape[0]
    out = torch.empty((T, 6, h, w), device=w2c.device, dtype=torch.float32)
    check(_lib.load().seva_plucker(w2c_src.data_ptr(), w2c.data_ptr(), Kpix.data_ptr(), out.data_ptr(), T, h, w,
                                   _stream()), "seva_plucker")
    return out


def resize_bilinear_ac(src: torch.Tensor, hd: int, wd: int, out: torch.Tensor | None = None) -> torch.Tensor:
    _req(src, torch.float32, "src")
    assert src.is_contiguous()
    N, Cc, hs, ws = src.shape
    if out is None:
        out = torch.empty((N, hd * wd, Cc), device=src.device, dtype=torch.float32)
    check(_lib.load().seva_resize_bilinear_ac(src.data_ptr(), out.data_ptr(), N, Cc, hs, ws, hd, wd, _stream()),
          "seva_resize_bilinear_ac")
    return out


def multiview_scale(c2w: torch.Tensor, K: torch.Tensor, input_mask: torch.Tensor, scale, min_scale: float):
    """Returns (scale[T] fp32, close_frame[T] bool)."""
    _req(c2w, torch.float32, "c2w"), _req(K, torch.float32, "K")
    T = c2w.shape[0]
    c2w4 = c2w.contiguous()
    assert c2w4.shape[-2:] == (4, 4), "c2w must be [T,4,4]"
    mask_u8 = input_mask.to(device=c2w.device, dtype=torch.uint8).contiguous()  # the reference accepts a CPU mask
    K = K.to(c2w.device)
    if isinstance(scale, torch.Tensor):
        sc = scale.reshape(-1).to(device=c2w.device, dtype=torch.float32).contiguous()
        is_vec = int(sc.numel() > 1)
        assert sc.numel() in (1, T)
    else:
        sc = torch.full((1,), float(scale), device=c2w.device, dtype=torch.float32)
        is_vec = 0
    out = torch.empty((T,), device=c2w.device, dtype=torch.float32)
    close = torch.empty((T,), device=c2w.device, dtype=torch.uint8)
    check(_lib.load().seva_multiview_scale(c2w4.data_ptr(), K.contiguous().data_ptr(), mask_u8.data_ptr(),
                                           sc.data_ptr(), is_vec, float(min_scale), out.data_ptr(),
                                           close.data_ptr(), T, _stream()), "seva_multiview_scale")
    return out, close.bool()


# ------------------------------------------------------------------------------------------------
# sampler-side
# ------------------------------------------------------------------------------------------------
def sigma_quantize(sigma: torch.Tensor, table: torch.Tensor):
    _req(sigma, torch.float32, "sigma"), _req(table, torch.float32, "table")
    N = sigma.numel()
    idx = torch.empty((N,), device=sigma.device, dtype=torch.int64)
    sq = torch.empty((N,), device=sigma.device, dtype=torch.float32)
    c_in = torch.empty_like(sq)
    c_out = torch.empty_like(sq)
    check(_lib.load().seva_sigma_quantize(sigma.contiguous().data_ptr(), table.data_ptr(), table.numel(),
                                          idx.data_ptr(), sq.data_ptr(), c_in.data_ptr(), c_out.data_ptr(), N,
                                          _stream()), "seva_sigma_quantize")
    return idx, sq, c_in, c_out


def _req_on(dev, dtype, **tensors) -> None:
    """Every tensor a kernel dereferences must live on the launch device with the dtype the kernel assumes."""
    for name, t in tensors.items():
        if t is None:
            continue
        _req(t, dtype, name)
        if t.device != dev:
            raise _lib.SevaB200Error(f"{name} is on {t.device}, expected {dev}")


def add_noise(x: torch.Tensor, eps: torch.Tensor, sigma: torch.Tensor, sigma_hat: torch.Tensor) -> torch.Tensor:
    _req_on(x.device, torch.float32, x=x, eps=eps, sigma=sigma, sigma_hat=sigma_hat)
    x, eps = x.contiguous(), eps.contiguous()
    out = torch.empty_like(x)
    T = x.shape[0]
    check(_lib.load().seva_add_noise(x.data_ptr(), eps.data_ptr(), sigma.contiguous().data_ptr(),
                                     sigma_hat.contiguous().data_ptr(), out.data_ptr(), T, x.numel() // T,
                                     _stream()), "seva_add_noise")
    return out


def denoise_prepare(x, replace, concat, c_in, net_in, blended_out, N, x_frames, h, w) -> None:
    _req_on(net_in.device, torch.float32, x=x, replace=replace, concat=concat, c_in=c_in, blended_out=blended_out)
    if tuple(concat.shape) != (N, 7, h, w) or (replace is not None and tuple(replace.shape) != (N, 5, h, w)) \
            or tuple(x.shape) != (x_frames, 4, h, w) or c_in.numel() != N or net_in.shape[1] < 16:
        raise _lib.SevaB200Error("seva_denoise_prepare: expected x [x_frames,4,h,w], replace [N,5,h,w], concat [N,7,h,w], "
                                 f"c_in [N]; got {tuple(x.shape)}, {None if replace is None else tuple(replace.shape)}, "
                                 f"{tuple(concat.shape)}, {tuple(c_in.shape)}")
    if not (x.is_contiguous() and concat.is_contiguous() and (replace is None or replace.is_contiguous())):
        raise _lib.SevaB200Error("seva_denoise_prepare: inputs must be contiguous")
    check(_lib.load().seva_denoise_prepare(x.data_ptr(), _p(replace), concat.data_ptr(), c_in.data_ptr(),
                                           net_in.data_ptr(), net_in.stride(0), blended_out.data_ptr(), N,
                                           x_frames, h, w, _stream()), "seva_denoise_prepare")


def denoise_finish(eps, blended, c_out, out, N, h, w) -> None:
    _req_on(eps.device, torch.float32, eps=eps, blended=blended, c_out=c_out, out=out)
    if eps.shape[1] < 4 or tuple(blended.shape) != (N, 4, h, w) or tuple(out.shape) != (N, 4, h, w) or c_out.numel() != N:
        raise _lib.SevaB200Error("seva_denoise_finish: expected eps [N*(h+2)*(w+2), >=4], blended / out [N,4,h,w], c_out [N]")
    check(_lib.load().seva_denoise_finish(eps.data_ptr(), eps.stride(0), blended.data_ptr(), c_out.data_ptr(),
                                          out.data_ptr(), N, h, w, _stream()), "seva_denoise_finish")


def cfg_euler_step(x: torch.Tensor, denoised: torch.Tensor, scale: torch.Tensor, sigma_hat: torch.Tensor,
                   sigma_next: torch.Tensor, out: torch.Tensor | None = None,
                   denoised_out: torch.Tensor | None = None) -> torch.Tensor:
    """x [T,...] fp32, denoised [2T,...] (uncond first), scale/sigma_hat/sigma_next [T] -> x_next."""
    _req_on(x.device, torch.float32, x=x, denoised=denoised, scale=scale, sigma_hat=sigma_hat, sigma_next=sigma_next,
            out=out, denoised_out=denoised_out)
    x, denoised = x.contiguous(), denoised.contiguous()
    T = x.shape[0]
    assert denoised.shape[0] == 2 * T and scale.numel() == T and sigma_hat.numel() == T and sigma_next.numel() == T
    if out is None:
        out = torch.empty_like(x)
    check(_lib.load().seva_cfg_euler_step(x.data_ptr(), denoised.data_ptr(), scale.data_ptr(),
                                          sigma_hat.data_ptr(), sigma_next.data_ptr(), out.data_ptr(),
                                          _p(denoised_out), T, x.numel() // T, _stream()), "seva_cfg_euler_step")
    return out


def pack_nchw(src: torch.Tensor, dst: torch.Tensor) -> None:
    _req(src, torch.float32, "src"), _req(dst, torch.bfloat16, "dst")
    N, Cc, h, w = src.shape
    check(_lib.load().seva_pack_nchw(src.contiguous().data_ptr(), dst.data_ptr(), dst.stride(0), N, Cc, h, w,
                                     _stream()), "seva_pack_nchw")


def unpack_nchw(src: torch.Tensor, N: int, Cc: int, h: int, w: int, out: torch.Tensor | None = None) -> torch.Tensor:
    _req(src, torch.float32, "src")
    if out is None:
        out = torch.empty((N, Cc, h, w), device=src.device, dtype=torch.float32)
    check(_lib.load().seva_unpack_nchw(src.data_ptr(), src.stride(0), out.data_ptr(), N, Cc, h, w, _stream()),
          "seva_unpack_nchw")
    return out


# ------------------------------------------------------------------------------------------------
# data movement / small dense
# ------------------------------------------------------------------------------------------------
def upsample2x_pad(src: torch.Tensor, dst: torch.Tensor, N: int, h: int, w: int, Cc: int) -> None:
    check(_lib.load().seva_upsample2x_pad(src.data_ptr(), src.stride(0), dst.data_ptr(), dst.stride(0), N, h, w, Cc,
                                          _stream()), "seva_upsample2x_pad")


def im2col3x3_s2(src: torch.Tensor, dst: torch.Tensor, N: int, h: int, w: int, Cc: int, origin: int = 0) -> None:
    """``origin`` 0: symmetric pad 1 (Seva Downsample); 1: pad right / bottom only (the VAE's Downsample2D)."""
    check(_lib.load().seva_im2col3x3_s2_at(src.data_ptr(), src.stride(0), dst.data_ptr(), N, h, w, Cc, origin,
                                           _stream()), "seva_im2col3x3_s2_at")


def softmax_rows(x: torch.Tensor, out: torch.Tensor | None = None) -> torch.Tensor:
    """fp32 scores [rows, L] -> bf16 probabilities [rows, L]."""
    _req(x, torch.float32, "x")
    rows, L = x.shape
    if out is None:
        out = torch.empty((rows, L), device=x.device, dtype=torch.bfloat16)
    check(_lib.load().seva_softmax_rows(x.data_ptr(), x.stride(0), out.data_ptr(), out.stride(0), rows, L, _stream()),
          "seva_softmax_rows")
    return out


def cast_bf16(src: torch.Tensor) -> torch.Tensor:
    _req(src, torch.float32, "src")
    src = src.contiguous()
    dst = torch.empty(src.shape, device=src.device, dtype=torch.bfloat16)
    check(_lib.load().seva_cast_f32_bf16(src.data_ptr(), dst.data_ptr(), src.numel(), _stream()),
          "seva_cast_f32_bf16")
    return dst


def timestep_embedding(t: torch.Tensor, dim: int, max_period: float = 10000.0,
                       out: torch.Tensor | None = None) -> torch.Tensor:
    _req(t, torch.int64, "t")
    N = t.numel()
    if out is None:
        out = torch.empty((N, dim), device=t.device, dtype=torch.float32)
    check(_lib.load().seva_timestep_embedding(t.contiguous().data_ptr(), out.data_ptr(), N, dim, max_period,
                                              _stream()), "seva_timestep_embedding")
    return out


def gemv_rows(x: torch.Tensor, W: torch.Tensor, bias: torch.Tensor | None, *, silu_in=False, silu_out=False,
              out: torch.Tensor | None = None) -> torch.Tensor:
    _req(x, torch.float32, "x"), _req(W, torch.bfloat16, "W")
    assert x.is_contiguous() and W.is_contiguous()
    M, K = x.shape
    N = W.shape[0]
    if out is None:
        out = torch.empty((M, N), device=x.device, dtype=torch.float32)
    check(_lib.load().seva_gemv_rows(x.data_ptr(), W.data_ptr(), _p(bias), out.data_ptr(), M, N, K, int(silu_in),
                                     int(silu_out), _stream()), "seva_gemv_rows")
    return out


# ------------------------------------------------------------------------------------------------
# CLIP conditioner / image preprocessing
# ------------------------------------------------------------------------------------------------
def resize_bicubic_aa(x: torch.Tensor, size: tuple[int, int], antialias: bool = True) -> torch.Tensor:
    """kornia.geometry.resize(x, size, interpolation="bicubic", align_corners=True, antialias=...) on fp32 [..., H, W]."""
    _req(x, torch.float32, "x")
    x = x.contiguous()
    H, W = x.shape[-2:]
    out = torch.empty((*x.shape[:-2], size[0], size[1]), device=x.device, dtype=torch.float32)
    check(_lib.load().seva_resize_bicubic_aa(x.data_ptr(), out.data_ptr(), x.numel() // (H * W), H, W, size[0], size[1],
                                             int(antialias), _stream()), "seva_resize_bicubic_aa")
    return out


def resize_area(x: torch.Tensor, size: tuple[int, int]) -> torch.Tensor:
    """F.interpolate(x, size, mode="area") on fp32 [..., H, W]."""
    _req(x, torch.float32, "x")
    x = x.contiguous()
    H, W = x.shape[-2:]
    out = torch.empty((*x.shape[:-2], size[0], size[1]), device=x.device, dtype=torch.float32)
    check(_lib.load().seva_resize_area(x.data_ptr(), out.data_ptr(), x.numel() // (H * W), H, W, size[0], size[1],
                                       _stream()), "seva_resize_area")
    return out


def clip_patchify(x: torch.Tensor, patch: int, rows_per_img: int, ldo: int, mean, std) -> torch.Tensor:
    """x fp32 [N, 3, S, S] in [-1, 1] -> normalised bf16 patch rows [N * rows_per_img, ldo] (row 0 of every image = 0)."""
    import ctypes

    _req(x, torch.float32, "x")
    x = x.contiguous()
    N, C, S, S2 = x.shape
    if C != 3 or S != S2:
        raise _lib.SevaB200Error(f"clip_patchify: expected [N, 3, S, S], got {tuple(x.shape)}")
    out = torch.empty((N * rows_per_img, ldo), device=x.device, dtype=torch.bfloat16)
    m = (ctypes.c_float * 3)(*[float(v) for v in mean])
    s = (ctypes.c_float * 3)(*[float(v) for v in std])
    check(_lib.load().seva_clip_patchify(x.data_ptr(), out.data_ptr(), N, S, patch, rows_per_img, ldo, m, s, _stream()),
          "seva_clip_patchify")
    return out
