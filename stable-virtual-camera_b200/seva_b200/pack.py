"""Weight packing into the layouts the kernels consume (bf16, K-major; see DESIGN.md "Data layout")."""
from __future__ import annotations

import torch


def pack_linear(w: torch.Tensor) -> torch.Tensor:
    """nn.Linear weight [N, K] -> bf16 [N, K] (already K-major)."""
    return w.detach().to(torch.bfloat16).contiguous()


def pack_conv3x3(w: torch.Tensor, cin_pad: int | None = None) -> torch.Tensor:
    """nn.Conv2d weight [Cout, Cin, 3, 3] -> bf16 [Cout, 9*Cin'] with column (ky*3+kx)*Cin' + c."""
    cout, cin, kh, kw = w.shape
    assert kh == 3 and kw == 3
    w = w.detach().permute(0, 2, 3, 1)  # [Cout, 3, 3, Cin]
    if cin_pad is not None and cin_pad > cin:
        w = torch.nn.functional.pad(w, (0, cin_pad - cin))
        cin = cin_pad
    return w.reshape(cout, 9 * cin).to(torch.bfloat16).contiguous()


def pack_conv1x1(w: torch.Tensor) -> torch.Tensor:
    """nn.Conv2d 1x1 weight [Cout, Cin, 1, 1] -> bf16 [Cout, Cin]."""
    return w.detach().reshape(w.shape[0], w.shape[1]).to(torch.bfloat16).contiguous()


def geglu_perm(inner: int, bn: int, device=None) -> torch.Tensor:
    """Row permutation that interleaves [x | gate] per N tile of width bn (see gemm.cu epilogue).

    Projection rows are [x_0..x_{I-1}, g_0..g_{I-1}] (GEGLU.forward chunk(2), transformer.py:14).  Packed tile t
    holds x rows [t*bn/2, (t+1)*bn/2) followed by the matching gate rows.
    """
    assert bn % 64 == 0 and (2 * inner) % bn == 0
    half = bn // 2
    idx = []
    for t in range(2 * inner // bn):
        idx.append(torch.arange(t * half, (t + 1) * half))
        idx.append(inner + torch.arange(t * half, (t + 1) * half))
    return torch.cat(idx).to(device) if device is not None else torch.cat(idx)


def pack_geglu(w: torch.Tensor, b: torch.Tensor, bn: int):
    inner = w.shape[0] // 2
    perm = geglu_perm(inner, bn, w.device)
    return w.detach()[perm].to(torch.bfloat16).contiguous(), b.detach()[perm].float().contiguous()


def pack_upconv_phases(w: torch.Tensor) -> list[torch.Tensor]:
    """nn.Conv2d weight [Cout, Cin, 3, 3] of Upsample.conv (layers.py:42-46) -> four bf16 [Cout, 4*Cin] matrices, one per
    output phase (py, px) in the order 2*py + px, columns (ty*2 + tx)*Cin + c.

    out[2y+py, 2x+px] of conv3x3(nearest2x(v)) only sees v[y+py-1 .. y+py, x+px-1 .. x+px]: the three kernel rows fall on two
    source rows -- py = 0: {w0} on y-1, {w1 + w2} on y;  py = 1: {w0 + w1} on y, {w2} on y+1 -- and likewise the columns.
    Sums are taken in fp32 before the bf16 rounding."""
    cout, cin, kh, kw = w.shape
    assert kh == 3 and kw == 3
    w = w.detach().float()
    rows = {0: (w[:, :, 0], w[:, :, 1] + w[:, :, 2]), 1: (w[:, :, 0] + w[:, :, 1], w[:, :, 2])}   # each [Cout, Cin, 3]
    out = []
    for py in (0, 1):
        for px in (0, 1):
            taps = []
            for ty in (0, 1):
                r = rows[py][ty]
                cols = (r[:, :, 0], r[:, :, 1] + r[:, :, 2]) if px == 0 else (r[:, :, 0] + r[:, :, 1], r[:, :, 2])
                taps += [cols[0], cols[1]]
            out.append(torch.cat(taps, dim=1).to(torch.bfloat16).contiguous())
    return out
