"""ORACLE — test infrastructure only.

Generates the golden vectors under tests/golden/ by executing the REAL reference (imported from
/root/reference, build container only) on the seeded synthetic inputs of oracle/synth.py:

    python oracle/gen_golden.py            # schedule, geometry, tiny-network vectors (seconds)
    python oracle/gen_golden.py --full     # + config 1: Seva 1.3B, T=8 @72x72, one Euler step, fp32 CPU (~2 min)

The reference has no tests or fixtures of its own; these files are the parity pin.
"""
from __future__ import annotations

import argparse
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import refimport, synth  # noqa: E402

GOLD = os.path.join(ROOT, "tests", "golden")


def save(name, **arrays):
    os.makedirs(GOLD, exist_ok=True)
    path = os.path.join(GOLD, name + ".npz")
    np.savez_compressed(path, **{k: (v.detach().cpu().numpy() if isinstance(v, torch.Tensor) else np.asarray(v))
                                 for k, v in arrays.items()})
    print(f"wrote {path} ({os.path.getsize(path) / 1e6:.2f} MB)")


def ref_plucker(ref):
    def fn(w2c_src, w2c, K, h, w):
        return ref.geometry.get_plucker_coordinates(extrinsics_src=w2c_src, extrinsics=w2c, intrinsics=K,
                                                    target_size=(h, w))
    return fn


def gen_schedule(ref):
    S = ref.sampling
    disc = S.DDPMDiscretization()
    den = S.DiscreteDenoiser(disc, num_idx=1000, device="cpu")
    out = {"table": den.sigmas}
    for n in (50, 25, 10, 3):
        sig = disc(n, device="cpu")
        out[f"sigmas_{n}"] = sig
        out[f"idx_{n}"] = den.sigma_to_idx(sig[:-1] + 1e-6)
    probe = torch.tensor([0.0, 0.0246, 0.03, 1.0, 14.5768, 84.9163, 100.0, 53.6834 + 1e-6])
    out["probe"], out["probe_idx"] = probe, den.sigma_to_idx(probe)
    out["probe_sigma"] = den.idx_to_sigma(den.sigma_to_idx(probe))
    save("schedule", **out)


def gen_geometry(ref):
    G, S = ref.geometry, ref.sampling
    out = {}
    for tag, (T, h, w) in {"a": (5, 24, 24), "b": (2, 72, 72), "c": (3, 16, 40)}.items():
        c2w = synth.orbit_cameras(T, radius=2.0 + 0.3 * T, arc=0.4)
        w2c = torch.linalg.inv(c2w)
        K = G.get_default_intrinsics().repeat(T, 1, 1)
        if tag == "c":
            K = G.get_default_intrinsics(aspect_ratio=w / h).repeat(T, 1, 1)
        out[f"w2c_{tag}"], out[f"K_{tag}"] = w2c, K
        out[f"plucker_{tag}"] = G.get_plucker_coordinates(w2c[0], w2c, K.clone(), target_size=(h, w))
    # guidance bookkeeping: duplicate poses / different intrinsics / small rotations
    T = 8
    c2w = synth.orbit_cameras(T, arc=0.5)
    K = G.get_default_intrinsics().repeat(T, 1, 1)
    mask = torch.zeros(T, dtype=torch.bool)
    mask[0] = mask[5] = True
    c2w[3] = c2w[0]                      # same pose as an input -> close
    c2w[6] = c2w[5]
    K[6, 0, 0] += 0.01                   # same pose, different K -> not close
    c2w[2, :3, 3] = c2w[0, :3, 3]        # same position, rotated by the orbit yaw (> 10 deg) -> not close
    c2w[7] = c2w[5]
    c2w[7, :3, 3] += 2e-5                # moved by more than 1e-5 -> not close
    rule = S.MultiviewScaleRule(min_scale=1.2)
    out["g_c2w"], out["g_K"], out["g_mask"] = c2w, K, mask
    out["g_rot"] = G.get_camera_dist(c2w, c2w[mask], mode="rotation")
    out["g_trans"] = G.get_camera_dist(c2w, c2w[mask], mode="translation")
    out["g_scale_float"] = rule(2.0, c2w, K, mask)
    out["g_scale_tensor"] = rule(torch.linspace(1.5, 3.0, T), c2w, K, mask)
    tcfg = S.MultiviewTemporalCFG(T, cfg_min=1.2)
    xu = torch.arange(2 * T * 4, dtype=torch.float32).reshape(2 * T, 4, 1, 1) / 10
    out["g_temporal_in"] = xu
    out["g_temporal_out"] = tcfg(xu, torch.ones(T), 3.0, c2w, K, mask)
    out["g_mv_out"] = S.MultiviewCFG(1.2)(xu, torch.ones(T), 2.0, c2w, K, mask)
    out["g_vanilla_out"] = S.VanillaCFG()(xu, torch.ones(T), 2.0)
    save("geometry", **out)


def run_reference_sampling(ref, model, inp, T, num_steps, scale, cfg_min, guider="multiview"):
    """Reference EulerEDMSampler loop with the pre-drawn eps injected through torch.randn_like."""
    S = ref.sampling
    disc = S.DDPMDiscretization()
    denoiser = S.DiscreteDenoiser(disc, num_idx=1000, device="cpu")
    g = {"multiview": lambda: S.MultiviewCFG(cfg_min), "temporal": lambda: S.MultiviewTemporalCFG(T, cfg_min),
         "vanilla": lambda: S.VanillaCFG()}[guider]()
    sampler = S.EulerEDMSampler(discretization=disc, guider=g, num_steps=num_steps, s_churn=0.0, s_tmin=0.0,
                                s_tmax=999.0, s_noise=1.0, verbose=False, device="cpu")
    wrapped = ref.model.SGMWrapper(model)
    captured = []

    def den_fn(input, sigma, c):
        out = denoiser(wrapped, input, sigma, c, num_frames=T)
        captured.append(out)
        return out

    eps_iter = iter(inp["eps"])
    orig = torch.randn_like
    torch.randn_like = lambda t, *a, **k: next(eps_iter).to(t)
    try:
        kw = {} if guider == "vanilla" else dict(c2w=inp["c2w"], K=inp["K"], input_frame_mask=inp["mask"])
        with torch.inference_mode():
            x = sampler(den_fn, inp["x0"].clone(), scale=scale, cond=dict(inp["c"]), uc=dict(inp["uc"]),
                        verbose=False, **kw)
    finally:
        torch.randn_like = orig
    return x, captured


def gen_tiny(ref):
    M = ref.model
    torch.manual_seed(0)
    params = M.SevaParams(model_channels=64, num_frames=3)
    model = M.Seva(params).eval()
    sd = synth.fill_state_dict({k: v.shape for k, v in model.state_dict().items()}, seed=0)
    model.load_state_dict(sd)
    T, h, w = 3, 16, 16
    inp = synth.make_inputs(T, 1, h, w, ref_plucker(ref), seed=1, num_steps=3, ctx_dim=params.context_dim)
    # one raw network call on a CFG batch
    g = torch.Generator().manual_seed(5)
    xin = torch.randn(2 * T, 4, h, w, generator=g)
    tt = torch.tensor([999, 999, 999, 500, 500, 500])
    cc = {k: torch.cat((inp["uc"][k], inp["c"][k]), 0) for k in ("crossattn", "concat", "dense_vector")}
    with torch.inference_mode():
        net_out = M.SGMWrapper(model)(xin, tt, cc, num_frames=T)
    x3, cap = run_reference_sampling(ref, model, inp, T, 3, 2.0, 1.2)
    xt, _ = run_reference_sampling(ref, model, inp, T, 3, 3.0, 1.2, guider="temporal")
    save("tiny", keys=np.array(sorted(sd)), plucker=inp["plucker"], net_in=xin, net_t=tt, net_out=net_out,
         x_final=x3, denoised_step0=cap[0], x_final_temporal=xt)


def gen_full(ref):
    M = ref.model
    t0 = time.time()
    params = M.SevaParams()
    with torch.device("meta"):
        model = M.Seva(params)
    shapes = {k: v.shape for k, v in model.state_dict().items()}
    sd = synth.fill_state_dict(shapes, seed=0)
    model.load_state_dict(sd, assign=True)
    model.eval()
    print(f"reference model built in {time.time() - t0:.1f}s, {sum(v.numel() for v in sd.values()) / 1e9:.3f} B params")
    T, h, w = 8, 72, 72
    inp = synth.make_inputs(T, 1, h, w, ref_plucker(ref), seed=1, num_steps=1)
    torch.set_num_threads(os.cpu_count())
    t0 = time.time()
    x1, cap = run_reference_sampling(ref, model, inp, T, 1, 2.0, 1.2)
    dt = time.time() - t0
    print(f"config 1 reference step: {dt:.1f}s on {torch.get_num_threads()} threads -> {T / dt:.4f} frame-steps/s")
    save("config1", plucker=inp["plucker"].half(), x1=x1, denoised=cap[0].half(), n_keys=len(sd),
         cpu_seconds=dt, cpu_threads=torch.get_num_threads())


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--full", action="store_true")
    ap.add_argument("--only-full", action="store_true")
    args = ap.parse_args()
    ref = refimport.import_reference()
    if not args.only_full:
        gen_schedule(ref)
        gen_geometry(ref)
        gen_tiny(ref)
    if args.full or args.only_full:
        gen_full(ref)
