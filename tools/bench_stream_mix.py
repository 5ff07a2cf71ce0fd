"""Developer tool: what the memory system gives for the residual linear's traffic pattern (two bf16 [M, 320] streams read,
one written) with no GEMM in the way -- the HBM floor the K = 320 output projections are measured against.
Usage: python tools/bench_stream_mix.py"""
import torch

M, N = 217728, 320
dev = "cuda"
a = torch.randn(M, N, device=dev).bfloat16()
b = torch.randn(M, N, device=dev).bfloat16()
c = torch.empty_like(a)
flush = torch.empty(256 * 1024 * 1024, device=dev, dtype=torch.uint8)


def timeit(fn, n=7):
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


nbytes = a.numel() * 2
for name, fn, streams in (("copy  (1 read, 1 write)", lambda: c.copy_(a), 2),
                          ("add   (2 reads, 1 write)", lambda: torch.add(a, b, out=c), 3)):
    t = timeit(fn)
    print(f"{name:26s} {t * 1e3:7.1f} us  {streams * nbytes / t / 1e9:6.2f} TB/s  ({streams * nbytes / 1e6:.0f} MB)")
