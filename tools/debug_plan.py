"""Developer tool: run one plan launch by launch and report the first launch whose outputs are non-finite."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "stable-virtual-camera_b200")]
import torch

import bench
from seva_b200 import _lib, ops

T = int(sys.argv[1]) if len(sys.argv) > 1 else 21
size = int(sys.argv[2]) if len(sys.argv) > 2 else 72
dev = torch.device("cuda", 0)
model, P = bench.build_model(dev)
host = bench.make_host_inputs(T, size, size, seed=1)
eng = model.module.engine(2 * T, T, size, size, 1)
x = torch.cat([host["x0"].cuda() * 84.9] * 2)
c = {k: torch.cat((host["uc"][k].cuda(), host["c"][k].cuda()), 0) for k in host["c"]}
sigma = torch.full((2 * T,), 84.9163, device=dev)
from seva_b200 import sampling as S
den = S.DiscreteDenoiser(S.DDPMDiscretization(), 1000, dev)
idx, _, c_in, c_out = ops.sigma_quantize(sigma, den.sigmas)
blended = torch.empty(2 * T, 4, size, size, device=dev)
ops.denoise_prepare(x, c["replace"], c["concat"], c_in, eng.net_in, blended, 2 * T, 2 * T, size, size)
eng.bind(idx, c["crossattn"], c["dense_vector"])
stream = torch.cuda.current_stream().cuda_stream
bad = set()
for i, (fn, args) in enumerate(eng.calls):
    rc = fn(*args, stream)
    assert rc == 0, _lib.last_error()
    torch.cuda.synchronize()
    for key, t in eng.bufs.items():
        if key in bad:
            continue
        if not torch.isfinite(t.float()).all():
            bad.add(key)
            nz = (~torch.isfinite(t.float())).nonzero()
            print(f"launch {i} {eng.meta[i][:2]} -> buffer {key} non-finite: {nz.shape[0]} elems, first {nz[0].tolist()}, absmax finite {t.float().nan_to_num(0,0,0).abs().max().item():.3e}")
    if len(bad) >= 3:
        break
print("eps finite:", torch.isfinite(eng.eps).all().item(), "absmax", eng.eps.abs().max().item())
