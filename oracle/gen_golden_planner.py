"""ORACLE — test infrastructure only.

Generates tests/golden/planner.json by running the REAL reference `seva.eval` bookkeeping functions (imported from
/root/reference, build container only) on the seeded cases of oracle/planner_cases.py:

    python oracle/gen_golden_planner.py

pad_indices, infer_prior_stats, infer_prior_inds, compute_relative_inds, chunk_input_and_test (every strategy family)
and one whole two-pass plan (1 input -> 80 targets, T=21).  Integer outputs only; the tests demand
equality.
"""
from __future__ import annotations

import copy
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import planner_cases as PC  # noqa: E402
from oracle import refimport  # noqa: E402


def tolist(x):
    if isinstance(x, np.ndarray):
        return x.tolist()
    if isinstance(x, (list, tuple)):
        return [tolist(v) for v in x]
    if isinstance(x, (np.integer,)):
        return int(x)
    if isinstance(x, (np.floating,)):
        return float(x)
    return x


def guarded(fn):
    try:
        return {"ok": tolist(fn())}
    except (AssertionError, NotImplementedError, IndexError, ValueError) as e:
        return {"error": type(e).__name__}


def main():
    E = refimport.import_eval()

    out = {}
    out["pad"] = [guarded(lambda c=c: E.pad_indices(list(c["input"]), list(c["test"]), c["T"], c["mode"]))
                  for c in PC.pad_cases()]

    def stats(c):
        v = {"T": copy.deepcopy(c["T"]), "options": dict(c["options"])}
        n = E.infer_prior_stats(copy.deepcopy(c["T"]), c["n_in"], c["n_total"], v)
        return [n, v["T"]]
    out["prior_stats"] = [guarded(lambda c=c: stats(c)) for c in PC.prior_stats_cases()]
    out["prior_inds"] = [guarded(lambda c=c: E.infer_prior_inds(PC.cameras(c["n"], 5), c["n_prior"], c["inputs"],
                                                                {"chunk_strategy": c["strategy"]}))
                         for c in PC.prior_inds_cases()]
    out["relative"] = [guarded(lambda c=c: [float(v) for v in E.compute_relative_inds(np.array(c["src"]), c["tgt"])])
                       for c in PC.relative_cases()]

    def chunks(c):
        ic, tc, io, to, gt = PC.chunk_inputs(c, E.infer_prior_stats, E.infer_prior_inds)
        opt = dict(c["options"], sampler_verbose=False)
        return E.chunk_input_and_test(c["T"], ic, tc, io, to, options=opt, task=c["task"], chunk_strategy=c["strategy"],
                                      gt_input_inds=list(gt))
    out["chunks"] = [guarded(lambda c=c: chunks(c)) for c in PC.chunk_cases()]
    n_err = sum("error" in r for r in out["chunks"])
    print(f"chunk cases: {len(out['chunks'])} ({n_err} raise in the reference)")

    # get_value_dict (eval.py:1152-1215): camera normalisation + ray map of one window, reference on CPU
    import torch
    vd = []
    for seed, T, far in ((11, 6, False), (12, 5, True)):
        allc = PC.cameras(30, seed)
        if far:
            allc[3, :3, 3] += 1e4  # an outlier camera: excluded from the centring mean by the quantile rule
        idx = list(range(0, 30, 30 // T))[:T]
        f = 0.5 / np.tan(0.5 * 0.9424777960769379)
        Ks = torch.tensor([[f, 0, 0.5], [0, f, 0.5], [0, 0, 1.0]], dtype=torch.float32).repeat(T, 1, 1)
        v = E.get_value_dict(torch.zeros(T, 3, 128, 128), torch.zeros(T, 3, 128, 128), [0], allc[idx].clone(), Ks,
                             list(range(T)), all_c2ws=allc, camera_scale=2.0)
        vd.append(dict(seed=seed, T=T, far=far, idx=idx, c2w=v["c2w"].tolist(),
                       plucker=v["plucker_coordinate"].tolist()))
    out["value_dict"] = vd

    path = os.path.join(ROOT, "tests", "golden", "planner.json")
    with open(path, "w") as f:
        json.dump(out, f, separators=(",", ":"))
    print(f"wrote {path} ({os.path.getsize(path) / 1e3:.1f} kB)")


if __name__ == "__main__":
    main()
