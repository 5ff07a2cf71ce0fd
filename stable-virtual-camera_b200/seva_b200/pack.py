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
