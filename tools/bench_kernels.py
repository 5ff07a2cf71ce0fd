"""Developer tool: time the HBM-bound kernels in isolation at the shapes of the T=21 step (CUDA events, L2 flushed)."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "stable-virtual-camera_b200")]
import torch
from seva_b200 import ops
dev = "cuda"
flush = torch.empty(256 * 1024 * 1024, device=dev, dtype=torch.uint8)
which = sys.argv[1] if len(sys.argv) > 1 else "all"

def timeit(fn, n=5):
    fn(); torch.cuda.synchronize()
    ts = []
    for _ in range(n):
        flush.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record(); torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    return min(ts)

N = 42
if which in ("all", "gn"):
    for (H, C, dense) in [(72, 320, True), (72, 320, False), (72, 640, True), (72, 960, True), (36, 640, True), (36, 1920, True), (18, 1280, True), (18, 2560, True), (9, 1280, True)]:
        rows = N * (H + 2) * (H + 2)
        x = torch.randn(rows, C, device=dev).bfloat16()
        y = torch.empty_like(x)
        gamma = torch.ones(C, device=dev); beta = torch.zeros(C, device=dev)
        dmap = torch.randn(N, H * H, 6, device=dev); dw = torch.randn(2 * C, 6, device=dev); db = torch.randn(2 * C, device=dev)
        sums = torch.zeros(ops.gn_stats_floats(N, (H + 2) * (H + 2), C), device=dev)
        t1 = timeit(lambda: ops.groupnorm_stats(x, N, (H + 2) * (H + 2), C, sums))
        p = ops.make_gn_apply(x_ptr=x.data_ptr(), ldx=C, y_ptr=y.data_ptr(), ldy=C, sums_ptr=sums.data_ptr(), gamma_ptr=gamma.data_ptr(), beta_ptr=beta.data_ptr(), n_img=N, H=H, Wd=H, Cc=C, groups=32, eps=1e-5, in_padded=True, out_padded=True, silu=True, dense_ptr=dmap.data_ptr() if dense else None, dense_w_ptr=dw.data_ptr() if dense else None, dense_b_ptr=db.data_ptr() if dense else None)
        t2 = timeit(lambda: ops.run_gn_apply(p))
        act = N * H * H * C * 2
        print(f"gn H={H} C={C} dense={dense}: stats {t1*1e3:.1f} us {act/t1/1e6:.0f} GB/s | apply {t2*1e3:.1f} us {2*act/t2/1e6:.0f} GB/s")
if which in ("all", "ln"):
    for (M, C) in [(217728, 320), (54432, 640), (13608, 1280)]:
        x = torch.randn(M, C, device=dev).bfloat16(); y = torch.empty_like(x)
        g = torch.ones(C, device=dev); b = torch.zeros(C, device=dev)
        t = timeit(lambda: ops.layernorm(x, g, b, out=y))
        print(f"ln M={M} C={C}: {t*1e3:.1f} us {4*M*C/t/1e6:.0f} GB/s")
if which in ("all", "tm"):
    for (S, heads) in [(5184, 5), (1296, 10), (324, 20)]:
        C = heads * 64
        qkv = torch.randn(2 * 21 * S, 3 * C, device=dev).bfloat16(); o = torch.empty(2 * 21 * S, C, device=dev, dtype=torch.bfloat16)
        t = timeit(lambda: ops.timemix_attention(qkv, 2, 21, S, heads, out=o))
        print(f"timemix S={S} heads={heads}: {t*1e3:.1f} us {8*2*21*S*C/t/1e6:.0f} GB/s")
if which in ("all", "gemv"):
    for (M, N_, K) in [(42, 1280, 1280), (42, 320, 1280), (42, 320, 1024), (2, 1280, 1024), (42, 320, 320)]:
        x = torch.randn(M, K, device=dev); W = torch.randn(N_, K, device=dev).bfloat16(); b = torch.zeros(N_, device=dev)
        o = torch.empty(M, N_, device=dev)
        t = timeit(lambda: ops.gemv_rows(x, W, b, silu_in=True, out=o))
        print(f"gemv M={M} N={N_} K={K}: {t*1e3:.1f} us")
if which in ("all", "step"):
    # the fused CFG-combine + Euler update moves ~10 MB at T=21 (latency-bound); the x64 batch (T=1344)
    # makes the HBM figure meaningful.  bytes = x, uncond, cond read + x_next written, fp32.
    for T in (21, 1344):
        x = torch.randn(T, 4, 72, 72, device=dev); den = torch.randn(2 * T, 4, 72, 72, device=dev)
        sc = torch.full((T,), 2.0, device=dev); s0 = torch.full((T,), 3.0, device=dev); s1 = torch.full((T,), 2.5, device=dev)
        out = torch.empty_like(x)
        t = timeit(lambda: ops.cfg_euler_step(x, den, sc, s0, s1, out=out))
        print(f"cfg_euler_step T={T}: {t*1e3:.1f} us {4*x.numel()*4/t/1e6:.0f} GB/s")
