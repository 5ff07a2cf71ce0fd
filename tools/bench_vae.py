"""Developer tool: time the VAE decode / encode (N3) at the benchmark's frame size on one B200 and put the same
network on library kernels (the oracle restatement under fp16 autocast, what seva/eval.py:1242 runs) beside it.
Usage: python tools/bench_vae.py [frames=21] [latent=72] [chunk=frames]"""
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "stable-virtual-camera_b200")]
import torch

from oracle import vae_oracle as VO
from seva_b200.autoencoder import AutoEncoder

T = int(sys.argv[1]) if len(sys.argv) > 1 else 21
hw = int(sys.argv[2]) if len(sys.argv) > 2 else 72
chunk = int(sys.argv[3]) if len(sys.argv) > 3 else T
dev = "cuda"
sd = {k: v.to(dev) for k, v in VO.random_state_dict(0).items()}
ae = AutoEncoder(state_dict=sd, seed=None, chunk_size=chunk).to(dev)
g = torch.Generator(device=dev).manual_seed(0)
z = torch.randn(T, 4, hw, hw, device=dev, generator=g) * 0.18215 * 4
x = torch.rand(2, 3, 8 * hw, 8 * hw, device=dev, generator=g) * 2 - 1


def timeit(fn, n=3):
    fn()
    torch.cuda.synchronize()
    ts = []
    for _ in range(n):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        fn()
        e1.record()
        torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    return min(ts)


t_dec = timeit(lambda: ae.decode(z))
t_enc = timeit(lambda: ae.encode(x))
print(f"ours: decode {T} x {8 * hw}^2 (chunk {chunk}): {t_dec:.1f} ms = {t_dec / T:.2f} ms/frame; encode 2 frames {t_enc:.1f} ms")
with torch.no_grad(), torch.autocast("cuda", dtype=torch.float16):
    t_ref = timeit(lambda: VO.decode(z, sd, chunk_size=1), n=2)
    t_ref_e = timeit(lambda: VO.encode(x, sd, chunk_size=1), n=2)
print(f"library kernels (restatement, fp16 autocast, decoding_t = 1): decode {t_ref:.1f} ms = {t_ref / T:.2f} ms/frame; "
      f"encode 2 frames {t_ref_e:.1f} ms")
with torch.no_grad():
    ref = VO.decode(z[:1], sd)
got = ae.decode(z[:1])
print(f"decode rel-L2 vs fp32 restatement (frame 0): {((got - ref).norm() / ref.norm()).item():.3e}")
print(f"peak memory {torch.cuda.max_memory_allocated() / 2**30:.1f} GiB")
