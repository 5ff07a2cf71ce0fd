"""ORACLE — test infrastructure only.

Seeded inputs for the chunk-planner parity fixtures (tests/golden/planner.json): the generator
(oracle/gen_golden_planner.py) feeds them to the REAL reference `seva.eval` functions, the tests feed the same
inputs to seva_b200.planner and compare bit for bit.  Everything is rebuilt from integers, so only the reference's
outputs are stored in the fixture.
"""
from __future__ import annotations

import math

import numpy as np
import torch


def cameras(n: int, seed: int) -> torch.Tensor:
    """[n,4,4] c2w along a wobbly arc: smooth translation plus seeded jitter, rotation about a seeded axis."""
    g = torch.Generator().manual_seed(seed)
    t = torch.linspace(0, 1, n)
    pos = torch.stack([2 * torch.sin(2.5 * t), 0.3 * torch.cos(4 * t), -2 * torch.cos(2.5 * t)], 1)
    pos = pos + 0.05 * torch.randn(n, 3, generator=g)
    axis = torch.nn.functional.normalize(torch.randn(3, generator=g), dim=0)
    ang = 1.5 * t + 0.05 * torch.randn(n, generator=g)
    Kx = torch.tensor([[0, -axis[2], axis[1]], [axis[2], 0, -axis[0]], [-axis[1], axis[0], 0]])
    R = torch.eye(3)[None] + torch.sin(ang)[:, None, None] * Kx + (1 - torch.cos(ang))[:, None, None] * (Kx @ Kx)
    c2w = torch.eye(4).repeat(n, 1, 1)
    c2w[:, :3, :3] = R
    c2w[:, :3, 3] = pos
    return c2w


def pad_cases() -> list:
    rng = np.random.RandomState(0)
    out = []
    for _ in range(40):
        T = int(rng.randint(3, 26))
        used = int(rng.randint(2, T + 1))
        slots = rng.permutation(T)[:used]
        n_in = int(rng.randint(1, used))
        out.append(dict(input=sorted(slots[:n_in].tolist()), test=sorted(slots[n_in:].tolist()), T=T,
                        mode="last" if rng.rand() < 0.8 else "none"))
    return out


def prior_stats_cases() -> list:
    out = []
    for strategy in ("nearest", "nearest-gt", "interp", "interp-gt", "gt"):
        for T in (21, [21, 21], [25, 21], 8):
            for n_in in (1, 2, 4, 9, 12):
                for n_total in (20, 80, 134):
                    for extra in ({}, {"num_prior_frames_ratio": 1.5}, {"num_prior_frames": 30}):
                        out.append(dict(T=T, n_in=n_in, n_total=n_total, options=dict(chunk_strategy=strategy, **extra)))
    return out


def prior_inds_cases() -> list:
    rng = np.random.RandomState(1)
    out = []
    for strategy in ("interp", "nearest"):
        for _ in range(12):
            n = int(rng.randint(12, 140))
            n_in = int(rng.randint(1, 5))
            inputs = sorted(rng.permutation(n)[:n_in].tolist())
            if rng.rand() < 0.5:
                inputs[0] = 0
                inputs = sorted(set(inputs))
            out.append(dict(n=n, n_prior=int(rng.randint(2, min(24, n - n_in))), inputs=inputs, strategy=strategy))
    return out


def relative_cases() -> list:
    rng = np.random.RandomState(2)
    out = []
    for _ in range(20):
        n = int(rng.randint(10, 100))
        src = sorted(rng.permutation(n)[: int(rng.randint(3, 9))].tolist())
        tgt = sorted(rng.permutation(n)[: int(rng.randint(1, 12))].tolist())
        out.append(dict(src=src, tgt=tgt))
    return out


def chunk_cases() -> list:
    """chunk_input_and_test calls.  `interp*` cases are built the way run_one_scene builds them (anchors from
    infer_prior_inds, orders = positions along the trajectory) so the reference's own assertions hold."""
    out = []
    seed = 100
    for strategy in ("gt", "gt-ltr", "gt-nearest"):
        for (T, M, N) in ((21, 1, 20), (21, 1, 47), (21, 3, 80), (8, 2, 19), (25, 4, 60)):
            for opt in ({}, {"pseudo_num_ratio": 0.5}, {"pseudo_num_max": 3}):
                seed += 1
                out.append(dict(strategy=strategy, T=T, M=M, N=N, options=opt, task="img2img", seed=seed,
                                gt=list(range(M)), layout="split"))
    for strategy in ("nearest", "nearest-gt", "nearest-1", "nearest-3", "nearest-5"):
        for (T, M, N) in ((21, 21, 80), (21, 9, 40), (8, 5, 33), (21, 30, 134), (21, 2, 5)):
            seed += 1
            n_gt = min(2, M)
            out.append(dict(strategy=strategy, T=T, M=M, N=N, options={}, task="img2trajvid", seed=seed,
                            gt=list(range(n_gt)), layout="split"))
    for strategy in ("interp", "interp-gt"):
        for task in ("img2trajvid", "img2vid"):
            for (T, n_in, n_total) in ((21, 1, 80), (21, 2, 134), (21, 1, 20), (25, 3, 60), (21, 1, 200)):
                seed += 1
                out.append(dict(strategy=strategy, T=T, M=None, N=n_total, options={}, task=task, seed=seed,
                                gt=list(range(n_in)), layout="traj", n_in=n_in))
    return out


def chunk_inputs(case: dict, infer_prior_stats, infer_prior_inds):
    """(input_c2ws, test_c2ws, input_ords, test_ords, gt_input_inds) of one chunk case.  The anchor count /
    positions of `traj` layouts come from the implementation under test (reference or ours): they are themselves
    covered by the prior_* fixtures."""
    if case["layout"] == "split":
        c = cameras(case["M"] + case["N"], case["seed"])
        perm = np.random.RandomState(case["seed"]).permutation(case["M"] + case["N"])
        in_idx, te_idx = sorted(perm[: case["M"]].tolist()), sorted(perm[case["M"]:].tolist())
        return c[in_idx], c[te_idx], in_idx, te_idx, case["gt"]
    # second pass of a two-pass run over a trajectory of N target frames (eval.py:1805-1850)
    n_in, n_total, T = case["n_in"], case["N"], case["T"]
    version = {"T": [T, T], "options": dict(chunk_strategy=case["strategy"], **case["options"])}
    n_prior = infer_prior_stats([T, T], n_in, n_total, version)
    traj = cameras(n_total, case["seed"])
    if case["task"] == "img2trajvid":
        inputs = cameras(n_in, case["seed"] + 7)
        input_indices = list(range(n_in))
        test_indices = list(range(n_in, n_in + n_total))
        prior = infer_prior_inds(traj, n_prior, [], version["options"])
        prior_indices = [int(i) + n_in for i in prior]
        all_c = torch.cat([inputs, traj])
    else:  # inputs are frames of the trajectory itself
        input_indices = sorted(np.random.RandomState(case["seed"]).permutation(n_total)[:n_in].tolist())
        test_indices = [i for i in range(n_total) if i not in input_indices]
        prior = infer_prior_inds(traj, n_prior, input_indices, version["options"])
        prior_indices = [int(i) for i in prior]
        all_c = traj
    order = np.argsort(input_indices + prior_indices).tolist()
    src_indices = np.array(input_indices + prior_indices)[order].tolist()
    gt = [order.index(i) for i in range(n_in)]
    return all_c[src_indices], all_c[test_indices], src_indices, test_indices, gt
