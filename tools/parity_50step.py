#!/usr/bin/env python
"""Parity of a full sampling run at BASELINE configs[1] size (Seva 1.3 B random-init, T=21 @576x576, 50 Euler-EDM steps,
MultiviewCFG) against the fp32 oracle restatement executed on the SAME GPU (eager PyTorch, TF32 off): final latents and
the denoiser output at steps {0, 25, 49} teacher-forced on our trajectory.  Test infrastructure (uses oracle/).

    python tools/parity_50step.py [--T 21] [--size 72] [--steps 50] > profiles/r1_parity_50step.json
"""
import argparse
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "stable-virtual-camera_b200")]
import torch

import bench
from oracle import seva_oracle as O
from oracle import synth
from seva_b200 import geometry as G
from seva_b200 import sampling as S

ap = argparse.ArgumentParser()
ap.add_argument("--T", type=int, default=21)
ap.add_argument("--size", type=int, default=72)
ap.add_argument("--steps", type=int, default=50)
a = ap.parse_args()
torch.backends.cuda.matmul.allow_tf32 = False
torch.backends.cudnn.allow_tf32 = False
dev = torch.device("cuda", 0)
T, h, w, steps = a.T, a.size, a.size, a.steps
model, P = bench.build_model(dev)
sd = {k: v.detach().float() for k, v in model.module.state_dict().items()}
plk = lambda s, e, K, hh, ww: G.get_plucker_coordinates(s, e, K, target_size=(hh, ww))
inp = synth.make_inputs(T, 1, h, w, plk, seed=1, num_steps=1)
cu = lambda d: {k: v.to(dev) for k, v in d.items()}
rel = lambda x, y: ((x.float() - y.float()).norm() / y.float().norm()).item()

disc = S.DDPMDiscretization()
denoiser = S.DiscreteDenoiser(disc, num_idx=1000, device=dev)
sampler = S.EulerEDMSampler(disc, S.MultiviewCFG(1.2), num_steps=steps, s_churn=0.0, s_tmin=0.0, s_tmax=999.0,
                            s_noise=1.0, verbose=False, device=dev)
traj, dens = [], []


def den_fn(x_in, sigma, c):
    traj.append((x_in.clone(), sigma.clone()))
    out = denoiser(model, x_in, sigma, c, num_frames=T)
    dens.append(out.clone())
    return out


eps_all = torch.randn(steps, T, 4, h, w, device=dev, generator=torch.Generator(device=dev).manual_seed(3))
eps_iter = iter(eps_all)
orig_randn_like = torch.randn_like
torch.randn_like = lambda t, *aa, **kk: next(eps_iter).to(t)   # the same noise draws for both implementations
t0 = time.time()
with torch.inference_mode(), torch.autocast("cuda"):
    ours = sampler(den_fn, inp["x0"].clone().to(dev), scale=2.0, cond=cu(inp["c"]), uc=cu(inp["uc"]), verbose=False,
                   c2w=inp["c2w"].to(dev), K=inp["K"].to(dev), input_frame_mask=inp["mask"].to(dev))
torch.cuda.synchronize()
t_ours = time.time() - t0
torch.randn_like = orig_randn_like

table = O.denoiser_table().to(dev)
fs = O.multiview_scale(2.0, 1.2, inp["c2w"], inp["K"], inp["mask"]).to(dev)
net = lambda xx, idx, c: O.sgm_forward(sd, P, xx, idx, c, num_frames=T)
c, uc = cu(inp["c"]), cu(inp["uc"])
t0 = time.time()
with torch.inference_mode():
    ref = O.sample(net, table, inp["x0"].to(dev), fs, c, uc, steps, eps_all)
    torch.cuda.synchronize()
    t_ref = time.time() - t0
    per_step = {}
    for i in sorted({0, steps // 2, steps - 1}):
        x_in, sigma = traj[i]                          # [2T,...] CFG batch our denoiser saw at step i
        d_ref = O.denoise(net, table, x_in, sigma, {k: torch.cat((uc[k], c[k]), 0) for k in c})
        per_step[str(i)] = rel(dens[i], d_ref)
# honest GPU baseline: the same restatement under bf16 autocast = eager PyTorch on library kernels
# (cuBLAS / cuDNN / flash SDPA), 3 steps after 1 warm-up
sig = torch.cat([O.ddpm_sigmas(steps), torch.zeros(1)]).to(dev)
ones = torch.ones(T, device=dev)
xb = inp["x0"].to(dev) * 84.9222
with torch.inference_mode(), torch.autocast("cuda", dtype=torch.bfloat16):
    for i in range(4):
        if i == 1:
            torch.cuda.synchronize()
            t0 = time.time()
        O.sampler_step(net, table, xb, ones * sig[i], ones * sig[i + 1], fs, c, uc, eps_all[i])
    torch.cuda.synchronize()
    t_bf16 = (time.time() - t0) / 3
print(json.dumps({
    "eager_pytorch_bf16_autocast_frame_steps_per_s": round(T / t_bf16, 2), "eager_pytorch_bf16_ms_per_step": round(t_bf16 * 1e3, 1),
    "what": f"Seva 1.3B random-init, T={T} @{h * 8}x{w * 8}, {steps} Euler-EDM steps, MultiviewCFG(1.2), scale 2.0; "
            "ours (bf16 kernels) vs oracle (fp32 eager PyTorch on the same B200, TF32 off)",
    "final_latents_rel_l2": rel(ours, ref), "gate_final": 5e-2,
    "denoiser_rel_l2_teacher_forced": per_step, "gate_step": 2e-2,
    "seconds_ours": round(t_ours, 2), "seconds_oracle_fp32_gpu": round(t_ref, 2),
    "oracle_frame_steps_per_s": round(T * steps / t_ref, 2), "gpu": torch.cuda.get_device_name(0)}))
