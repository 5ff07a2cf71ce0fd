"""Scene loop on the GPU: the window ray map vs the reference's get_value_dict fixture, and a small two-pass scene
(anchors, then chunked targets) through the CUDA path vs the same schedule sampled with the fp32 oracle."""
import json
import os

import numpy as np
import pytest
import torch

from oracle import planner_cases as PC
from oracle import seva_oracle as O
from oracle import synth
from seva_b200 import planner as P
from seva_b200 import sampling as S
from seva_b200 import scene as SC
from seva_b200.model import Seva, SevaParams, SGMWrapper

pytestmark = pytest.mark.gpu
GOLD = os.path.join(os.path.dirname(__file__), "golden", "planner.json")


def _K(n):
    f = 0.5 / np.tan(0.5 * 0.9424777960769379)
    return torch.tensor([[f, 0, 0.5], [0, f, 0.5], [0, 0, 1.0]], dtype=torch.float32).repeat(n, 1, 1)


def test_window_cameras_match_reference_value_dict():
    with open(GOLD) as f:
        cases = json.load(f)["value_dict"]
    for c in cases:
        allc = PC.cameras(30, c["seed"])
        if c["far"]:
            allc[3, :3, 3] += 1e4
        cam = SC.window_cameras(allc[c["idx"]], _K(c["T"]), allc, (16, 16), 2.0)
        want = torch.tensor(c["plucker"])
        assert cam["plucker"].shape == want.shape
        assert (cam["plucker"].cpu() - want).abs().max().item() < 2e-4
        assert torch.allclose(cam["c2w"].cpu(), torch.tensor(c["c2w"]), atol=1e-5, rtol=1e-5)


def test_two_pass_scene_matches_oracle_schedule():
    T, h, steps, n_targets = 6, 16, 3, 10
    prm = SevaParams(model_channels=64, num_frames=T)
    with torch.device("meta"):
        m = Seva(prm)
    sd = synth.fill_state_dict({k: tuple(v.shape) for k, v in m.state_dict().items()}, seed=0)
    m.load_state_dict(sd, assign=True)
    model = SGMWrapper(m.eval()).to("cuda")
    disc = S.DDPMDiscretization()
    denoiser = S.DiscreteDenoiser(disc, num_idx=1000, device="cuda")

    def make_sampler(T_, wrap):
        g = S.MultiviewCFG(1.2)
        return S.EulerEDMSampler(disc, wrap(g) if wrap else g, num_steps=steps, s_churn=0.0, s_tmin=0.0, s_tmax=999.0,
                                 s_noise=1.0, verbose=False, device="cuda")

    g = torch.Generator().manual_seed(5)
    traj, inp = PC.cameras(n_targets, 3), PC.cameras(1, 4)
    opts = {"chunk_strategy": "interp", "chunk_strategy_first_pass": "gt-nearest", "cfg": 2.0}
    version = {"T": [T, T], "options": opts}
    n_prior = P.infer_prior_stats([T, T], 1, n_targets, version)
    prior = P.infer_prior_inds(traj, n_prior, [], opts)
    sc = SC.Scene(torch.randn(1, 4, h, h, generator=g), inp, _K(1), traj, _K(n_targets), [0],
                  list(range(1, n_targets + 1)), traj[prior], _K(len(prior)), [int(i) + 1 for i in prior])
    plan = SC.plan_scene(sc, version["T"], opts, task="img2trajvid", use_traj_prior=True)
    assert len(plan.passes) == 2 and len(plan.passes[1].windows) >= 2
    codec = SC.Codec(dim=prm.context_dim)

    ours = SC.run_scene(sc, plan, model, denoiser, make_sampler, codec=codec,
                        generator=torch.Generator().manual_seed(7))
    torch.cuda.synchronize()

    table = O.denoiser_table()

    def oracle_sample(model_, den_, sampler_, win, rows, all_c2ws, x0, cfg, codec_, camera_scale, device):
        mask = torch.from_numpy(win.mask)
        Tw = len(mask)
        c2w, w2c = SC.normalize_cameras(rows["c2w"], all_c2ws, camera_scale)
        pl = O.plucker(w2c[0], w2c, rows["K"].clone(), h, h)
        c, uc = SC.window_cond(rows["img"][mask], mask, codec_.tokens(rows["img"][mask]).mean(0), pl)
        fs = O.multiview_scale(cfg, 1.2, c2w, rows["K"], mask)
        with torch.inference_mode():
            z = O.sample(lambda xx, idx, cc: O.sgm_forward(sd, prm, xx, idx, cc, num_frames=Tw), table, x0, fs, c, uc,
                         steps, torch.zeros(steps, *x0.shape))
        return z.clone()[win.test_sels]

    ref = SC.run_scene(sc, plan, None, None, lambda T_, wrap: None, codec=codec,
                       generator=torch.Generator().manual_seed(7), device="cpu", sample_fn=oracle_sample)
    assert ours.shape == ref.shape == (n_targets, 4, h, h)
    err = ((ours - ref).norm() / ref.norm()).item()
    assert err < 5e-2, err  # final latents after two chained bf16 passes vs fp32 (north_star: <= 5e-2)


def test_window_latents_do_not_depend_on_who_samples_them():
    """with the REAL sampler (per-step ``randn_like`` noise) a window must give the same latents whether it
    is sampled after the other windows on one rank or alone on another rank.  The scene seeds every window's noise
    stream from (scene seed, pass, window): sampling the second pass's windows in REVERSE order on a polluted RNG
    reproduces the sequential run bit for bit."""
    T, h, steps, n_targets = 4, 16, 2, 9
    prm = SevaParams(model_channels=64, num_frames=T)
    with torch.device("meta"):
        m = Seva(prm)
    sd = synth.fill_state_dict({k: tuple(v.shape) for k, v in m.state_dict().items()}, seed=0)
    m.load_state_dict(sd, assign=True)
    model = SGMWrapper(m.eval()).to("cuda")
    disc = S.DDPMDiscretization()
    denoiser = S.DiscreteDenoiser(disc, num_idx=1000, device="cuda")
    make_sampler = lambda T_, wrap: S.EulerEDMSampler(disc, S.MultiviewCFG(1.2), num_steps=steps, s_churn=0.0, s_tmin=0.0,
                                                      s_tmax=999.0, s_noise=1.0, verbose=False, device="cuda")
    g = torch.Generator().manual_seed(5)
    traj, inp = PC.cameras(n_targets, 3), PC.cameras(1, 4)
    opts = {"chunk_strategy": "interp", "chunk_strategy_first_pass": "gt-nearest", "cfg": 2.0}
    version = {"T": [T, T], "options": opts}
    prior = P.infer_prior_inds(traj, P.infer_prior_stats([T, T], 1, n_targets, version), [], opts)
    sc = SC.Scene(torch.randn(1, 4, h, h, generator=g), inp, _K(1), traj, _K(n_targets), [0],
                  list(range(1, n_targets + 1)), traj[prior], _K(len(prior)), [int(i) + 1 for i in prior])
    plan = SC.plan_scene(sc, version["T"], opts, task="img2trajvid", use_traj_prior=True)
    codec = SC.Codec(dim=prm.context_dim)
    seen = {}

    def recording(model_, den_, sampler_, win, rows, all_c2ws, x0, cfg, codec_, camera_scale, device):
        out = SC.sample_window(model_, den_, sampler_, win, rows, all_c2ws, x0, cfg, codec_, camera_scale, device)
        seen.setdefault("calls", []).append((win, rows, all_c2ws, x0, cfg, out))
        return out

    a = SC.run_scene(sc, plan, model, denoiser, make_sampler, codec=codec, generator=torch.Generator().manual_seed(7),
                     sample_fn=recording)
    b = SC.run_scene(sc, plan, model, denoiser, make_sampler, codec=codec, generator=torch.Generator().manual_seed(7))
    assert torch.equal(a, b)
    n_first = len(plan.passes[0].windows)
    second = seen["calls"][n_first:]
    assert len(second) >= 2
    sampler = make_sampler(T, None)
    torch.cuda.manual_seed(999)
    torch.randn(1000, device="cuda")  # another rank's RNG history
    for k in reversed(range(len(second))):
        win, rows, all_c2ws, x0, cfg, want = second[k]
        SC._seed_step_noise("cuda", SC.window_noise_seed(7, 1, k))
        got = SC.sample_window(model, denoiser, sampler, win, rows, all_c2ws, x0, cfg, codec, 2.0, "cuda")
        assert torch.equal(got, want), k


def test_scene_with_vae_and_clip_modules():
    """N3 + N4 either side of the path: a one-pass scene on RGB frames through AutoEncoder.encode, the CLIP conditioner,
    the sampler and AutoEncoder.decode (do_sample, eval.py:1245-1248, 1316) -- the whole per-window sequence on our kernels."""
    from seva_b200.autoencoder import AutoEncoder
    from seva_b200.conditioner import CLIPConditioner, ViTParams

    T, H, steps, n_targets = 4, 128, 2, 5
    prm = SevaParams(model_channels=64, num_frames=T)
    with torch.device("meta"):
        m = Seva(prm)
    sd = synth.fill_state_dict({k: tuple(v.shape) for k, v in m.state_dict().items()}, seed=0)
    m.load_state_dict(sd, assign=True)
    model = SGMWrapper(m.eval()).to("cuda")
    disc = S.DDPMDiscretization()
    denoiser = S.DiscreteDenoiser(disc, num_idx=1000, device="cuda")
    ae = AutoEncoder(chunk_size=None, seed=2).cuda()
    clip = CLIPConditioner(ViTParams(image_size=56, patch_size=14, width=128, layers=2, heads=2, embed_dim=prm.context_dim),
                           seed=3).cuda()
    codec = SC.Codec.from_modules(ae, clip)
    assert codec.latent_shape((3, H, H)) == (4, H // 8, H // 8)

    def make_sampler(T_, wrap):
        g = S.MultiviewCFG(1.2)
        return S.EulerEDMSampler(disc, wrap(g) if wrap else g, num_steps=steps, s_churn=0.0, s_tmin=0.0, s_tmax=999.0,
                                 s_noise=1.0, verbose=False, device="cuda")

    g = torch.Generator().manual_seed(5)
    traj, inp = PC.cameras(n_targets, 3), PC.cameras(1, 4)
    opts = {"chunk_strategy": "nearest-gt", "cfg": 2.0}
    frames = torch.rand(1, 3, H, H, generator=g) * 2 - 1
    sc = SC.Scene(frames, inp, _K(1), traj, _K(n_targets), [0], list(range(1, n_targets + 1)))
    plan = SC.plan_scene(sc, T, opts, task="img2trajvid", use_traj_prior=False)
    out = SC.run_scene(sc, plan, model, denoiser, make_sampler, codec=codec, generator=torch.Generator().manual_seed(7))
    assert out.shape == (n_targets, 3, H, H) and torch.isfinite(out).all()
    # the same run again returns the same frames (per-window noise streams, no hidden state in the codec modules)
    again = SC.run_scene(sc, plan, model, denoiser, make_sampler, codec=codec, generator=torch.Generator().manual_seed(7))
    assert torch.equal(out, again)
