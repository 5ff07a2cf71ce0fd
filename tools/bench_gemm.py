"""Developer tool: time seva_gemm_bf16 in isolation at the dominant shapes of the T=21 step (CUDA events, L2 flushed
between runs) and check each result against torch.  Usage: python tools/bench_gemm.py [filter]"""
import math
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "stable-virtual-camera_b200")]
import torch
import torch.nn.functional as F

from seva_b200 import ops, pack

dev = "cuda"
flush = torch.empty(256 * 1024 * 1024, device=dev, dtype=torch.uint8)
filt = sys.argv[1] if len(sys.argv) > 1 else ""
CL = int(os.environ.get("SEVA_CLUSTER", "0"))
BN = int(os.environ.get("SEVA_BN", "0"))
LNF = int(os.environ.get("SEVA_LNFOLD", "0"))   # 1: the geglu / no-residual linear cases fold a LayerNorm of A (row statistics given)
LNR = int(os.environ.get("SEVA_LNROWS", "0"))   # 1: the residual linear cases also write LayerNorm row statistics of their output


def timeit(fn, n=5):
    fn()
    torch.cuda.synchronize()
    ts = []
    for _ in range(n):
        flush.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        fn()
        e1.record()
        torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    return min(ts)


def rel(a, b):
    return ((a.float() - b.float()).norm() / b.float().norm()).item()


g = torch.Generator(device=dev).manual_seed(0)
rnd = lambda *s: torch.randn(*s, device=dev, generator=g)

# (kind, M, N, K, residual)
LIN = [("linear", 217728, 320, 320, True), ("linear", 217728, 320, 320, False), ("linear", 217728, 960, 320, False),
       ("linear", 217728, 320, 1280, True), ("linear", 54432, 640, 2560, True), ("linear", 54432, 640, 640, True),
       ("linear", 54432, 1920, 640, False), ("linear", 13608, 1280, 5120, True), ("linear", 13608, 1280, 1280, True),
       ("linear", 13608, 3840, 1280, False), ("linear", 3402, 1280, 1280, True),
       ("geglu", 217728, 2560, 320, False), ("geglu", 54432, 5120, 640, False), ("geglu", 13608, 10240, 1280, False)]
for kind, M, N, K, has_res in LIN:
    name = f"{kind} M{M} N{N} K{K} res{int(has_res)}"
    if filt and filt not in name:
        continue
    A = rnd(M, K).bfloat16()
    W = (rnd(N, K) / math.sqrt(K)).bfloat16()
    b = rnd(N) * 0.1
    ln_kw = {}
    if LNF and (kind == "geglu" or not has_res):
        # row statistics of A as its producer would have left them (K columns, split over the producer's slots)
        slots = ops.ln_slots(K)
        Af = A.float()
        st = torch.stack([Af.sum(1), (Af * Af).sum(1)], -1) / slots
        ln_kw = dict(ln_in=st[None].repeat(slots, 1, 1).contiguous(), ln_colsum=W.float().sum(1).contiguous())
    if kind == "geglu":
        bn = ops.pick_bn(N, True)
        Wp, bp = pack.pack_geglu(W, b, bn)
        if ln_kw:
            ln_kw["ln_colsum"] = pack.pack_geglu(W, W.float().sum(1), bn)[1]
        out = torch.empty(M, N // 2, device=dev, dtype=torch.bfloat16)
        fn = lambda: ops.linear(A, Wp, bp, geglu=True, out=out, cluster=CL, **ln_kw)
    else:
        res = rnd(M, N).bfloat16() if has_res else None
        out = torch.empty(M, N, device=dev, dtype=torch.bfloat16)
        rows_kw = dict(ln_rowstats=torch.empty(ops.ln_slots(N, BN), M, 2, device=dev)) if LNR and N in (320, 640, 1280) else {}
        fn = lambda: ops.linear(A, W, b, out=out, residual=res, cluster=CL, bn=BN, **ln_kw, **rows_kw)
    t = timeit(fn)
    def ref(rows):
        Ar = F.layer_norm(A[rows].float(), (K,), None, None, 1e-5) if ln_kw else A[rows].float()
        y = Ar @ W.float().t() + b
        if kind == "geglu":
            x_, g_ = y.chunk(2, -1)
            return x_ * F.gelu(g_)
        return y + res[rows].float() if has_res else y

    err = max(rel(out[:4096], ref(slice(0, 4096))), rel(out[-2048:], ref(slice(M - 2048, M))))
    print(f"{name:44s} {t * 1e3:8.1f} us  {2.0 * M * N * K / t / 1e9:7.1f} TF/s  relL2 {err:.2e}", flush=True)

# conv3x3: (n_img, H, Cin, Cout, residual)
CONV = [(42, 72, 320, 320, True), (42, 72, 640, 320, False), (42, 36, 640, 640, True), (42, 18, 1280, 1280, True),
        (42, 9, 1280, 1280, True), (42, 18, 2560, 1280, False)]
for n_img, H, Cin, Cout, has_res in CONV:
    name = f"conv3x3 n{n_img} H{H} Cin{Cin} Cout{Cout} res{int(has_res)}"
    if filt and filt not in name:
        continue
    Hp = H + 2
    ximg = torch.zeros(n_img, Hp, Hp, Cin, device=dev)
    ximg[:, 1:-1, 1:-1] = rnd(n_img, H, H, Cin)
    x = ximg.reshape(-1, Cin).bfloat16()
    Wc = (rnd(Cout, Cin, 3, 3) / math.sqrt(9 * Cin)).bfloat16()
    W9 = pack.pack_conv3x3(Wc)
    b = rnd(Cout) * 0.1
    res = None
    if has_res:
        rimg = torch.zeros(n_img, Hp, Hp, Cout, device=dev)
        rimg[:, 1:-1, 1:-1] = rnd(n_img, H, H, Cout)
        res = rimg.reshape(-1, Cout).bfloat16()
    out = torch.empty(n_img * Hp * Hp, Cout, device=dev, dtype=torch.bfloat16)
    fn = lambda: ops.conv3x3_padded(x, W9, b, n_img, H, H, out=out, residual=res, cluster=CL, bn=BN)
    t = timeit(fn)
    xi = x.reshape(n_img, Hp, Hp, Cin)[:2, 1:-1, 1:-1].permute(0, 3, 1, 2).float()
    y = F.conv2d(xi, Wc.float(), b, padding=1).permute(0, 2, 3, 1)
    if has_res:
        y = y + res.reshape(n_img, Hp, Hp, Cout)[:2, 1:-1, 1:-1].float()
    o = out.reshape(n_img, Hp, Hp, Cout)
    err = rel(o[:2, 1:-1, 1:-1], y)
    border = max(o[:, 0].abs().max().item(), o[:, -1].abs().max().item(), o[:, :, 0].abs().max().item(),
                 o[:, :, -1].abs().max().item())
    print(f"{name:44s} {t * 1e3:8.1f} us  {2.0 * n_img * H * H * Cout * Cin * 9 / t / 1e9:7.1f} TF/s  relL2 {err:.2e} "
          f"border {border:.1e}", flush=True)

# Upsample.forward as four phase convolutions (SevaGemm ntaps = 4, out_map = 2..5): (n_img, low-res H, Cin, Cout)
UPCONV = [(42, 36, 640, 640), (42, 18, 1280, 1280), (42, 9, 1280, 1280)]
for n_img, H, Cin, Cout in UPCONV:
    name = f"upconv n{n_img} H{H}->{2 * H} Cin{Cin} Cout{Cout}"
    if filt and filt not in name:
        continue
    Hp = H + 2
    ximg = torch.zeros(n_img, Hp, Hp, Cin, device=dev)
    ximg[:, 1:-1, 1:-1] = rnd(n_img, H, H, Cin)
    x = ximg.reshape(-1, Cin).bfloat16()
    Wc = rnd(Cout, Cin, 3, 3) / math.sqrt(9 * Cin)
    phases = pack.pack_upconv_phases(Wc)
    b = rnd(Cout) * 0.1
    out = torch.zeros(n_img * (2 * H + 2) * (2 * H + 2), Cout, device=dev, dtype=torch.bfloat16)
    gs = [ops.make_gemm(A_ptr=x.data_ptr(), lda=Cin, a_rows=x.shape[0], W_ptr=phases[ph].data_ptr(), ldw=4 * Cin,
                        M=x.shape[0], N=Cout, Kc=Cin, out_ptr=out.data_ptr(), ldo=Cout, ntaps=4, tap_wp=H + 2,
                        bias_ptr=b.data_ptr(), out_map=2 + ph, H=H, Wd=H) for ph in range(4)]

    def fn():
        for g_ in gs:
            ops.run_gemm(g_)

    t = timeit(fn)
    xi = x.reshape(n_img, Hp, Hp, Cin)[:2, 1:-1, 1:-1].permute(0, 3, 1, 2).float()
    y = F.conv2d(F.interpolate(xi, scale_factor=2, mode="nearest"), Wc, b, padding=1).permute(0, 2, 3, 1)
    o = out.reshape(n_img, 2 * H + 2, 2 * H + 2, Cout)
    err = rel(o[:2, 1:-1, 1:-1], y)
    print(f"{name:44s} {t * 1e3:8.1f} us  {2.0 * n_img * 4 * H * H * Cout * Cin * 9 / t / 1e9:7.1f} TF/s of reference FLOPs "
          f"relL2 {err:.2e}", flush=True)
