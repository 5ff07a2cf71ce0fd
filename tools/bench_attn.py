"""Developer tool: time seva_attn_fwd in isolation at the SDPA shapes of the T=21 step (CUDA events, L2
flushed between runs) and check against torch SDPA.  Usage: python tools/bench_attn.py [filter] [--once]"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "stable-virtual-camera_b200")]
import torch
import torch.nn.functional as F

from seva_b200 import ops

dev = "cuda"
args = [a for a in sys.argv[1:] if not a.startswith("--")]
once = "--once" in sys.argv
filt = args[0] if args else ""
flush = torch.empty(256 * 1024 * 1024, device=dev, dtype=torch.uint8)
g = torch.Generator(device=dev).manual_seed(0)

# (name, B, heads, L)
SHAPES = [("3d_ds2", 2, 10, 27216), ("3d_ds4", 2, 20, 6804), ("3d_ds8", 2, 20, 1701), ("2d_ds1", 42, 5, 5184),
          ("2d_ds2", 42, 10, 1296), ("2d_ds4", 42, 20, 324)]
for name, B, Hh, L in SHAPES:
    if filt and filt not in name:
        continue
    C = Hh * 64
    qkv = torch.randn(B * L, 3 * C, device=dev, generator=g).bfloat16()
    out = torch.empty(B * L, C, device=dev, dtype=torch.bfloat16)
    fn = lambda: ops.attention_qkv(qkv, B, L, Hh, out=out)
    fn()
    torch.cuda.synchronize()
    if once:
        continue
    ts = []
    for _ in range(5):
        flush.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        fn()
        e1.record()
        torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    t = min(ts)
    fl = 4.0 * B * Hh * L * L * 64
    q, k, v = (qkv.view(B, L, 3, Hh, 64)[:, :, i].transpose(1, 2) for i in range(3))
    nq = min(L, 2048)   # reference on a slice of the queries (full keys)
    ref = F.scaled_dot_product_attention(q[:, :, :nq].float(), k.float(), v.float()).transpose(1, 2).reshape(B, nq, C)
    got = out.view(B, L, C)[:, :nq].float()
    err = ((got - ref).norm() / ref.norm()).item()
    print(f"attn {name} B{B} H{Hh} L{L}: {t * 1e3:8.1f} us  {fl / t / 1e9:7.1f} TF/s  relL2 {err:.2e}", flush=True)
