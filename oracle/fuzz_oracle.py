"""ORACLE — test infrastructure only (build container: needs /root/reference).

Differential fuzz of the oracle's sampler-side restatement against the REAL reference (`seva.sampling`,
`seva.geometry`) on seeded random inputs, beyond the committed golden vectors: sigma tables for every step count,
sigma -> index quantisation (bit-exact), the multi-view / temporal guidance scale rules (bit-exact booleans,
including cameras placed on the 10 degree / 1e-5 thresholds), camera distances and Plucker ray maps (fp32, <= 1e-5).

    python oracle/fuzz_oracle.py [n=300] [seed=0]
"""
from __future__ import annotations

import math
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT]
from oracle import planner_cases as PC  # noqa: E402
from oracle import refimport  # noqa: E402
from oracle import seva_oracle as O  # noqa: E402


def rot(axis: torch.Tensor, deg: float) -> torch.Tensor:
    a = torch.nn.functional.normalize(axis, dim=0)
    K = torch.tensor([[0, -a[2], a[1]], [a[2], 0, -a[0]], [-a[1], a[0], 0]])
    t = math.radians(deg)
    return torch.eye(3) + math.sin(t) * K + (1 - math.cos(t)) * (K @ K)


def threshold_window(g: torch.Generator, i: int):
    """(T, c2w [T,4,4], K [T,3,3], input mask [T]) with some targets sharing an input pose up to a rotation on either
    side of the 10 degree rule, a translation on either side of 1e-5, or a changed focal length (sampling.py:171-180)."""
    T = int(torch.randint(2, 26, (1,), generator=g))
    c2w = PC.cameras(T, 5000 + i)
    K = O.default_intrinsics().repeat(T, 1, 1)
    mask = torch.rand(T, generator=g) < 0.3
    mask[int(torch.randint(0, T, (1,), generator=g))] = True
    src = torch.nonzero(mask).flatten()
    for t in range(T):
        if not mask[t] and torch.rand(1, generator=g) < 0.4:
            j = src[int(torch.randint(0, len(src), (1,), generator=g))]
            deg = float(torch.tensor([0.0, 5.0, 9.99, 10.01, 15.0])[int(torch.randint(0, 5, (1,), generator=g))])
            c2w[t] = c2w[j]
            c2w[t, :3, :3] = rot(torch.randn(3, generator=g), deg) @ c2w[j, :3, :3]
            if torch.rand(1, generator=g) < 0.3:
                c2w[t, 0, 3] += float(torch.tensor([5e-6, 2e-5])[int(torch.randint(0, 2, (1,), generator=g))])
            if torch.rand(1, generator=g) < 0.2:
                K[t, 0, 0] *= 1.01
    return T, c2w, K, mask


def main() -> int:
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 300
    g = torch.Generator().manual_seed(int(sys.argv[2]) if len(sys.argv) > 2 else 0)
    R = refimport.import_reference()
    S, G = R.sampling, R.geometry
    bad = 0

    # sigma tables, sampling.py:57-102
    disc = S.DDPMDiscretization()
    worst = 0.0
    for steps in list(range(1, 130)) + [200, 250, 500, 999, 1000]:
        want = disc(steps, do_append_zero=False, flip=False)
        got = O.ddpm_sigmas(steps)
        worst = max(worst, ((want - got).abs() / want.abs()).max().item())
    bad += worst > 1e-6
    print(f"ddpm sigmas: 134 step counts, max rel diff {worst:.2e}")

    # sigma -> idx, sampling.py:126-131 (bit-exact; includes exact table entries and midpoints)
    den = S.DiscreteDenoiser(disc, num_idx=1000, device="cpu")
    table = O.denoiser_table()
    assert torch.equal(table, den.sigmas)
    mism = 0
    for _ in range(n):
        s = torch.exp(torch.rand(64, generator=g) * 9.0 - 4.0)
        k = torch.randint(0, 999, (16,), generator=g)
        s = torch.cat([s, table[k], 0.5 * (table[k] + table[k + 1]), table[k] * (1 + 1e-7)])
        mism += int((den.sigma_to_idx(s) != O.sigma_to_idx(table, s)).sum())
    bad += mism > 0
    print(f"sigma_to_idx: {n * 112} sigmas, {mism} index mismatches")

    # guidance scale rules, sampling.py:160-213, 245-298
    mism = tmism = 0
    for i in range(n):
        T, c2w, K, mask = threshold_window(g, i)
        want = S.MultiviewScaleRule(min_scale=1.2)(2.0, c2w, K, mask)
        got = O.multiview_scale(2.0, 1.2, c2w, K, mask)
        mism += int((torch.as_tensor(want).float().expand(T) != torch.as_tensor(got).float().expand(T)).sum())
        guider = S.MultiviewTemporalCFG(T, 1.2)
        x = torch.randn(2 * T, 1, 2, 2, generator=g)
        want = guider(x.clone(), torch.ones(T), 3.0, c2w, K, mask)  # sigma_hat is [T] (sampling.py:363)
        sc = O.multiview_scale(O.temporal_scale(3.0, 1.2, mask, T), 1.2, c2w, K, mask)
        u, c = x.chunk(2)
        got = u + sc.view(-1, 1, 1, 1) * (c - u)
        tmism += int(not torch.allclose(want, got, rtol=1e-6, atol=1e-6))
    bad += (mism > 0) + (tmism > 0)
    print(f"multiview scale rule: {n} windows, {mism} frames differ; temporal CFG: {tmism} windows differ")

    # camera distances and ray maps, geometry.py:12-40, 119-165
    worst_d = worst_p = 0.0
    for i in range(max(n // 10, 1)):
        T = int(torch.randint(2, 22, (1,), generator=g))
        c2w = PC.cameras(T, 7000 + i)
        for mode in ("rotation", "translation"):
            worst_d = max(worst_d, (G.get_camera_dist(c2w, c2w[:3], mode=mode) - O.camera_dist(c2w, c2w[:3], mode)).abs().max().item())
        w2c = torch.linalg.inv(c2w)
        h, w = (int(v) for v in torch.randint(4, 40, (2,), generator=g))
        K = O.default_intrinsics().repeat(T, 1, 1)
        want = G.get_plucker_coordinates(w2c[0], w2c, intrinsics=K.clone(), target_size=[h, w])
        got = O.plucker(w2c[0], w2c, K, h, w)
        worst_p = max(worst_p, (want - got).abs().max().item())
    bad += (worst_d > 1e-3) + (worst_p > 1e-5)
    print(f"camera_dist: max abs diff {worst_d:.2e} (degrees / units); plucker: max abs diff {worst_p:.2e}")
    # network restatement vs the reference modules on several small configurations (model.py:176-234)
    from oracle import synth
    worst_n = 0.0
    for mc, T, h, w, heads_ch, depth in ((32, 2, 8, 8, 32, [1, 1, 1, 1]), (32, 3, 16, 24, 32, [1, 1, 1, 1]),
                                         (64, 5, 24, 16, 64, [1, 1, 1, 1]), (64, 4, 16, 16, 32, [1, 1, 1, 1]),
                                         (64, 3, 16, 16, 64, [2, 1, 3, 1])):
        P = R.model.SevaParams(model_channels=mc, num_frames=T, num_head_channels=heads_ch, transformer_depth=depth)
        torch.manual_seed(mc + T)
        net = R.model.SGMWrapper(R.model.Seva(P).eval())
        sd = {k: v.detach().clone() for k, v in net.module.state_dict().items()}
        inp = synth.make_inputs(T, 1, h, w, O.plucker, seed=T, num_steps=1, ctx_dim=P.context_dim)
        cc = {k: torch.cat((inp["uc"][k], inp["c"][k]), 0) for k in ("crossattn", "concat", "dense_vector")}
        x = torch.randn(2 * T, 4, h, w, generator=g)
        t = torch.randint(0, 1000, (1,), generator=g).repeat(2 * T)
        with torch.inference_mode():
            want = net(x, t, cc, num_frames=T)
            got = O.sgm_forward(sd, P, x, t, cc, num_frames=T)
        worst_n = max(worst_n, ((want - got).norm() / want.norm()).item())
    bad += worst_n > 1e-5
    print(f"network forward (5 small configurations incl. non-square latents and transformer_depth 2 / 3): max rel-L2 {worst_n:.2e}")
    return int(bad > 0)


if __name__ == "__main__":
    sys.exit(main())
