"""ORACLE — test infrastructure only (build container: needs /root/reference).

Differential fuzz of the chunk planner: random scenes are fed to the REAL reference `seva.eval` functions and to
`seva_b200.planner`; every output (lists of index lists, or the exception type) must be equal.  The committed fixture
(tests/golden/planner.json) pins 90 chunk scenes; this widens the same comparison to thousands of seeded scenes
without growing the fixture.  Prints one summary line per function family and exits non-zero on the first mismatch.

    python oracle/fuzz_planner.py [n_scenes=1500] [seed=0]
"""
from __future__ import annotations

import copy
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "stable-virtual-camera_b200")]
from oracle import planner_cases as PC  # noqa: E402
from oracle import refimport  # noqa: E402
from oracle.gen_golden_planner import tolist  # noqa: E402

def guarded(fn):
    """Result, or the exception type: the reference's failure modes (asserts, empty-range divisions) are part of
    the behaviour a drop-in has to reproduce."""
    try:
        return {"ok": tolist(fn())}
    except Exception as e:  # noqa: BLE001
        return {"error": type(e).__name__}


SPLIT = ("gt", "gt-ltr", "gt-nearest", "nearest", "nearest-gt", "nearest-1", "nearest-2", "nearest-3", "nearest-5")
TRAJ = ("interp", "interp-gt")


def scene(rng: np.random.RandomState, i: int) -> dict:
    T = int(rng.choice([8, 12, 21, 21, 21, 25]))
    if rng.rand() < 0.6:
        strategy = str(rng.choice(SPLIT))
        M = int(rng.randint(1, 34))
        N = int(rng.randint(1, 150))
        opt = {}
        if strategy.startswith("gt") and rng.rand() < 0.5:
            opt = {"pseudo_num_ratio": float(rng.choice([0.25, 0.5, 0.75]))} if rng.rand() < 0.5 else \
                  {"pseudo_num_max": int(rng.randint(1, 6))}
        n_gt = int(rng.randint(1, min(M, 3) + 1))
        return dict(strategy=strategy, T=T, M=M, N=N, options=opt, seed=1000 + i, layout="split",
                    task=str(rng.choice(["img2img", "img2trajvid"])), gt=list(range(n_gt)))
    n_in = int(rng.randint(1, 5))
    return dict(strategy=str(rng.choice(TRAJ)), T=T, M=None, N=int(rng.randint(max(T, 12), 260)), options={},
                task=str(rng.choice(["img2trajvid", "img2vid"])), seed=1000 + i, gt=list(range(n_in)), layout="traj",
                n_in=n_in)


def main() -> int:
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 1500
    rng = np.random.RandomState(int(sys.argv[2]) if len(sys.argv) > 2 else 0)
    E = refimport.import_eval()
    from seva_b200 import planner as P

    def chunks(mod, c):
        ic, tc, io, to, gt = PC.chunk_inputs(c, mod.infer_prior_stats, mod.infer_prior_inds)
        return mod.chunk_input_and_test(c["T"], ic, tc, io, to, options=dict(c["options"], sampler_verbose=False),
                                        task=c["task"], chunk_strategy=c["strategy"], gt_input_inds=list(gt))

    bad = raised = 0
    for i in range(n):
        c = scene(rng, i)
        want, got = guarded(lambda: chunks(E, c)), guarded(lambda: chunks(P, c))
        raised += "error" in want
        if want != got:
            bad += 1
            print("MISMATCH chunk", c, "\n  ref ", str(want)[:300], "\n  ours", str(got)[:300])
            break
    print(f"chunk_input_and_test: {n} scenes ({raised} raise in the reference, same exception type in ours), "
          f"{bad} mismatches")

    pbad = 0
    for i in range(n):
        T = int(rng.randint(3, 70))
        used = int(rng.randint(2, T + 1))
        slots = rng.permutation(T)[:used]
        n_in = int(rng.randint(1, used))
        a = (sorted(slots[:n_in].tolist()), sorted(slots[n_in:].tolist()), T, "last" if rng.rand() < 0.8 else "none")
        want = guarded(lambda: E.pad_indices(list(a[0]), list(a[1]), a[2], a[3]))
        got = guarded(lambda: P.pad_indices(list(a[0]), list(a[1]), a[2], a[3]))
        pbad += want != got
    print(f"pad_indices: {n} cases, {pbad} mismatches")

    sbad = 0
    for i in range(n):
        T = [int(rng.choice([8, 21, 25, 64]))] * 2 if rng.rand() < 0.7 else int(rng.choice([8, 21, 25]))
        c = dict(T=T, n_in=int(rng.randint(1, 30)), n_total=int(rng.randint(1, 400)),
                 options=dict(chunk_strategy=str(rng.choice(["nearest", "nearest-gt", "interp", "interp-gt", "gt"])),
                              **({} if rng.rand() < 0.6 else {"num_prior_frames_ratio": float(rng.choice([0.5, 1.5, 2.0]))})))

        def stats(mod):
            v = {"T": copy.deepcopy(c["T"]), "options": dict(c["options"])}
            return [mod.infer_prior_stats(copy.deepcopy(c["T"]), c["n_in"], c["n_total"], v), v["T"]]
        sbad += guarded(lambda: stats(E)) != guarded(lambda: stats(P))
    print(f"infer_prior_stats: {n} cases, {sbad} mismatches")

    ibad = 0
    for i in range(max(n // 5, 1)):
        m = int(rng.randint(12, 300))
        n_in = int(rng.randint(0, 5))
        inputs = sorted(rng.permutation(m)[:n_in].tolist())
        c = dict(n=m, n_prior=int(rng.randint(2, min(40, m - n_in))), inputs=inputs,
                 strategy=str(rng.choice(["interp", "nearest"])))
        cams = PC.cameras(m, 2000 + i)
        f = lambda mod: tolist(mod.infer_prior_inds(cams, c["n_prior"], list(c["inputs"]), {"chunk_strategy": c["strategy"]}))
        ibad += guarded(lambda: f(E)) != guarded(lambda: f(P))
    print(f"infer_prior_inds: {max(n // 5, 1)} cases, {ibad} mismatches")
    return int(bool(bad or pbad or sbad or ibad))


if __name__ == "__main__":
    sys.exit(main())
