import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "stable-virtual-camera_b200")]
import torch, bench
from seva_b200 import sampling as S, ops, _lib
T, h, w = 21, 72, 72
dev = torch.device("cuda", 0)
model, _ = bench.build_model(dev)
host = bench.make_host_inputs(T, h, w, seed=1)
disc = S.DDPMDiscretization()
denoiser = S.DiscreteDenoiser(disc, num_idx=1000, device=dev)
sampler = S.EulerEDMSampler(disc, S.MultiviewCFG(1.2), num_steps=50, s_churn=0.0, s_tmin=0.0, s_tmax=999.0, s_noise=1.0, verbose=False, device=dev)
sigmas = disc(50, device=dev)
den_fn = lambda i, s, c: denoiser(model, i, s, c, num_frames=T)
with torch.inference_mode(), torch.autocast("cuda"):
    nb = lambda t: t.to(dev, non_blocking=True)
    x = nb(host["x0"]) * 84.9222
    c = {k: nb(v) for k, v in host["c"].items()}; uc = {k: nb(v) for k, v in host["uc"].items()}
    c2w, K, mask = nb(host["c2w"]), nb(host["K"]), nb(host["mask"])
    ones = x.new_ones([T])
    for i in range(2):
        x = sampler.sampler_step(ones * sigmas[i], ones * sigmas[i + 1], den_fn, x, 2.0, c, uc, 0.0, c2w=c2w, K=K, input_frame_mask=mask)
    print("x after 2 steps finite", torch.isfinite(x).all().item())
    # step 2 by hand
    eng = next(iter(model.module._engines.values()))
    xin, sig, cond = sampler.guider.prepare_inputs(x, ones * sigmas[2] + 1e-6, c, uc)
    idx, _, c_in, c_out = ops.sigma_quantize(sig, denoiser.sigmas)
    print("idx", idx[:3].tolist(), "c_in", c_in[:2].tolist())
    blended = torch.empty(2 * T, 4, h, w, device=dev)
    ops.denoise_prepare(xin, cond["replace"], cond["concat"], c_in, eng.net_in, blended, 2 * T, 2 * T, h, w)
    eng.bind(idx, cond["crossattn"], cond["dense_vector"])
    stream = torch.cuda.current_stream().cuda_stream
    bad = set()
    for i, (fn, args) in enumerate(eng.calls):
        rc = fn(*args, stream)
        assert rc == 0, _lib.last_error()
        torch.cuda.synchronize()
        if i % 1 == 0:
            for key, t in eng.bufs.items():
                if key in bad: continue
                if not torch.isfinite(t.float()).all():
                    bad.add(key)
                    nz = (~torch.isfinite(t.float())).nonzero()
                    print(f"launch {i} {eng.meta[i][:2]} -> buffer {key} non-finite: {nz.shape[0]} elems, first {nz[0].tolist()}, last {nz[-1].tolist()} absmax finite {t.float().nan_to_num(0,0,0).abs().max().item():.3e}", flush=True)
        if len(bad) >= 2: break
    print("done; eps finite", torch.isfinite(eng.eps).all().item())
