"""Scene loop (seva_b200.scene) host logic on CPU: camera normalisation vs the reference's get_value_dict fixture,
conditioning assembly, the config-3 plan, and the multi-rank schedule over a world_size-2 gloo group (a stand-in
sampler replaces the CUDA path; the same code runs over NCCL on the GPUs)."""
import json
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from oracle import planner_cases as PC
from oracle import synth
from seva_b200 import planner as P
from seva_b200 import scene as SC

GOLD = os.path.join(os.path.dirname(__file__), "golden", "planner.json")


def _K(n):
    f = 0.5 / np.tan(0.5 * 0.9424777960769379)
    return torch.tensor([[f, 0, 0.5], [0, f, 0.5], [0, 0, 1.0]], dtype=torch.float32).repeat(n, 1, 1)


def test_normalize_cameras_matches_reference_value_dict():
    with open(GOLD) as f:
        cases = json.load(f)["value_dict"]
    for c in cases:
        allc = PC.cameras(30, c["seed"])
        if c["far"]:
            allc[3, :3, 3] += 1e4
        c2w, w2c = SC.normalize_cameras(allc[c["idx"]], allc, 2.0)
        want = torch.tensor(c["c2w"])
        assert torch.allclose(c2w, want, atol=1e-5, rtol=1e-5)
        assert abs(torch.norm(c2w[0, :3, 3]).item() - 2.0) < 1e-4
        assert torch.allclose(w2c @ c2w, torch.eye(4).expand_as(c2w), atol=1e-4)


def test_window_cond_matches_reference_layout():
    """Same arrangement as eval.py:1249-1281 (restated independently in oracle/synth.py, which the golden
    generator feeds to the reference)."""
    from oracle import seva_oracle as O

    inp = synth.make_inputs(5, 2, 8, 8, O.plucker, seed=4, ctx_dim=32)
    g = torch.Generator().manual_seed(4)
    lat = torch.randn(2, 4, 8, 8, generator=g)
    tok = torch.randn(32, generator=g)
    c, uc = SC.window_cond(lat, inp["mask"], tok, inp["plucker"])
    for k in ("crossattn", "replace", "concat", "dense_vector"):
        assert torch.equal(c[k], inp["c"][k]), k
        assert torch.equal(uc[k], inp["uc"][k]), k


def _scene_1_to_n(n_targets=80, h=8):
    """Config 3 shape: one input view, n targets along a trajectory, anchors from the planner."""
    g = torch.Generator().manual_seed(9)
    traj = PC.cameras(n_targets, 3)
    inp = PC.cameras(1, 4)
    T = [21, 21]
    opts = {"chunk_strategy": "interp", "chunk_strategy_first_pass": "gt-nearest", "cfg": [3.0, 2.0]}
    version = {"T": T, "options": opts}
    n_prior = P.infer_prior_stats(T, 1, n_targets, version)
    prior = P.infer_prior_inds(traj, n_prior, [], opts)
    sc = SC.Scene(input_latents=torch.randn(1, 4, h, h, generator=g), input_c2ws=inp, input_Ks=_K(1),
                  target_c2ws=traj, target_Ks=_K(n_targets), input_indices=[0],
                  target_indices=list(range(1, n_targets + 1)), prior_c2ws=traj[prior], prior_Ks=_K(len(prior)),
                  prior_indices=[int(i) + 1 for i in prior])
    return sc, version["T"], opts


def test_plan_scene_config3():
    sc, T, opts = _scene_1_to_n()
    plan = SC.plan_scene(sc, T, opts, task="img2trajvid", use_traj_prior=True)
    first, second = plan.passes
    assert len(first.windows) == 1 and not first.chained and first.cfg == 3.0
    assert len(second.windows) == 7 and not second.chained and second.cfg == 2.0
    assert [len(w.test_inds) for w in second.windows] == [13, 12, 13, 12, 13, 12, 5]
    assert plan.gt_input_inds == [0] and plan.prior_argsort == list(range(21))
    assert first.windows[0].mask.tolist() == [True] + [False] * 20
    # every second-pass window is a full T=21 forward: padded slots repeat the last selection
    for w in second.windows:
        assert ((w.input_maps != -1) ^ (w.test_maps != -1)).all() and len(w.input_maps) == 21


def _fake_sample(model, denoiser, sampler, win, rows, all_c2ws, x0, cfg, codec, camera_scale, device):
    """Deterministic stand-in for one sampling run: depends on the conditioning frames, their slots, the cameras, the
    noise and the guidance scale, so a wrong table row / order / noise stream cannot cancel out."""
    mask = torch.from_numpy(win.mask)
    c2w, _ = SC.normalize_cameras(rows["c2w"], all_c2ws, camera_scale)
    ctx = rows["img"][mask].mean(0, keepdim=True)
    pos = c2w[:, :3, 3].sum(-1).view(-1, 1, 1, 1)
    z = 0.5 * ctx + 0.1 * x0 + 0.01 * pos * cfg + 0.001 * torch.arange(len(mask)).view(-1, 1, 1, 1)
    z[mask] = rows["img"][mask]
    return z[win.test_sels]


def _run(sc, T, opts, where, one_pass=False):
    if one_pass:
        sc = SC.Scene(sc.input_latents, sc.input_c2ws, sc.input_Ks, sc.target_c2ws[:50], sc.target_Ks[:50], [0],
                      list(range(1, 51)))
        plan = SC.plan_scene(sc, 21, {"chunk_strategy": "gt-nearest"}, task="img2img", use_traj_prior=False)
        assert plan.passes[0].chained
    else:
        plan = SC.plan_scene(sc, T, opts, task="img2trajvid", use_traj_prior=True)
    g = torch.Generator().manual_seed(77)
    return SC.run_scene(sc, plan, None, None, lambda T_, wrap: None, where=where, generator=g, device="cpu",
                        sample_fn=_fake_sample)


def test_run_scene_single_rank_shapes_and_order():
    sc, T, opts = _scene_1_to_n()
    out = _run(sc, T, opts, SC.Rank())
    assert out.shape == (80, 4, 8, 8) and torch.isfinite(out).all()
    chained = _run(sc, T, opts, SC.Rank(), one_pass=True)
    assert chained.shape == (50, 4, 8, 8)


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, ret):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        sc, T, opts = _scene_1_to_n()
        where = SC.Rank(rank=rank, world=world)
        ret[rank] = dict(two_pass=_run(sc, T, opts, where), one_pass=_run(sc, T, opts, where, one_pass=True))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world", [2, 3])
def test_run_scene_two_ranks_equals_one_rank(world):
    """7 second-pass windows over 2 ranks (4 + 3) and 3 ranks (3 + 2 + 2): every rank ends with the one-rank result."""
    sc, T, opts = _scene_1_to_n()
    want2, want1 = _run(sc, T, opts, SC.Rank()), _run(sc, T, opts, SC.Rank(), one_pass=True)
    ctx = mp.get_context("spawn")
    with ctx.Manager() as mgr:
        ret = mgr.dict()
        port = _free_port()
        procs = [ctx.Process(target=_worker, args=(r, world, port, ret)) for r in range(world)]
        for p in procs:
            p.start()
        for p in procs:
            p.join(600)
            assert p.exitcode == 0, f"worker exit code {p.exitcode}"
        for r in range(world):
            assert torch.equal(ret[r]["two_pass"], want2), f"rank {r}: chunk-parallel second pass differs"
            assert torch.equal(ret[r]["one_pass"], want1), f"rank {r}: chained one-pass differs"


def test_prepared_conditioning_cache_host_logic():
    """SURVEY 8f N2 host side: the stacked conditioning is built once per (c, uc) and sampling loop, tagged with an
    epoch, rebuilt when a tensor is replaced or (where versions are tracked) written in place, and dropped by
    prepare_sampling_loop; the per-step dict is always a fresh one (the denoiser pops 'replace' from it)."""
    import torch

    from seva_b200 import sampling as S

    T = 3
    c = {"crossattn": torch.randn(T, 1, 8), "replace": torch.randn(T, 5, 4, 4), "concat": torch.randn(T, 7, 4, 4),
         "dense_vector": torch.randn(T, 6, 4, 4)}
    uc = {k: torch.zeros_like(v) for k, v in c.items()}
    g = S.MultiviewCFG(1.2)
    x, s = torch.randn(T, 4, 4, 4), torch.ones(T)
    xa, sa, a = g.prepare_inputs(x, s, c, uc)
    xb, sb, b = g.prepare_inputs(x, s, c, uc)
    assert xa.shape[0] == 2 * T and sa.shape[0] == 2 * T
    assert a is not b and a[S.COND_EPOCH] == b[S.COND_EPOCH]
    for k in c:
        assert a[k] is b[k] and torch.equal(a[k], torch.cat((uc[k], c[k]), 0))  # uncond first (sampling.py:236)
    a.pop("replace")
    assert "replace" in g.prepare_inputs(x, s, c, uc)[2]
    c["crossattn"].mul_(2.0)  # in-place write: version counter moves
    d = g.prepare_inputs(x, s, c, uc)[2]
    assert d[S.COND_EPOCH] != a[S.COND_EPOCH] and torch.equal(d["crossattn"][T:], c["crossattn"])
    c2 = dict(c, concat=c["concat"].clone())  # replaced tensor object
    e = g.prepare_inputs(x, s, c2, uc)[2]
    assert e[S.COND_EPOCH] != d[S.COND_EPOCH]
    with torch.inference_mode():  # inference tensors track no version: identity only, reset between loops
        ci = {k: v.clone() for k, v in c.items()}
        f1 = g.prepare_inputs(x, s, ci, uc)[2]
        f2 = g.prepare_inputs(x, s, ci, uc)[2]
        assert f1[S.COND_EPOCH] == f2[S.COND_EPOCH]
        sampler = S.EulerEDMSampler(S.DDPMDiscretization(), g, num_steps=2, device="cpu")
        sampler.prepare_sampling_loop(x.clone(), ci, uc)
        assert g.prepare_inputs(x, s, ci, uc)[2][S.COND_EPOCH] != f1[S.COND_EPOCH]
    # non-tensor keys pass through and must agree between c and uc (sampling.py:238-239)
    c3, uc3 = dict(c, tag=1), dict(uc, tag=1)
    assert g.prepare_inputs(x, s, c3, uc3)[2]["tag"] == 1
    with pytest.raises(AssertionError):
        g.prepare_inputs(x, s, c3, dict(uc, tag=2))
