#!/usr/bin/env python
"""Benchmark of the Seva denoising step (BASELINE.json metric: frame-steps/s, T=21 @576x576).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--T 21] [--size 72]

ours       one "step" = one Euler-EDM sampler step with MultiviewCFG (CFG batch 2T) through the public
           seva_b200.sampling / seva_b200.model API on our sm_100a kernels, 1.3 B-parameter Seva, random-init
           weights, synthetic inputs.  `value` times K steps with everything resident in HBM; `e2e` repeats the
           K steps with the step's inputs copied from pinned host memory and the updated latents read back
           inside the timed region.  N > 1: one process per GPU, each rank samples its own independent T-frame
           chunk (the two-pass procedure's second-pass chunks) - no data-path collective;
           the timed region is bracketed by barrier + synchronize and the max over ranks is used (weak scaling).
           The line also carries `cpu_baseline` (one step of the reference's own code on the host cores) and
           `gpu_library_baseline` (the unmodified reference modules in eager PyTorch, bf16 autocast, on the same GPU).
reference  the REFERENCE's own implementation of the path (oracle/_ref, staged by oracle/make_ref.py: the
           unmodified seva/{model,sampling,geometry}.py + seva/modules/{layers,transformer}.py) on the box's host
           cores, fp32, all threads, at the FULL workload (T=21, CFG batch 42).  A step costs minutes, so the run is
           bounded by --cpu-budget seconds (at least one timed step; the line reports the steps actually timed).
           Falls back to the oracle port (and says so in cpu_baseline.kind) only when oracle/_ref is absent.
Prints ONE JSON line.
"""
from __future__ import annotations

import argparse
import json
import math
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
for _p in (ROOT, os.path.join(ROOT, "stable-virtual-camera_b200")):
    if _p not in sys.path:
        sys.path.insert(0, _p)

import torch  # noqa: E402

# Algorithmic FLOPs of the reference graph per CFG'd step (FlopCounterMode)
REF_TFLOP_PER_STEP = {8: 26.336, 21: 76.912, 64: 314.534}


def measured_peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        with open(path) as f:
            d = json.load(f)
        return dict(hbm=d["hbm_gbs"], tf_burst=d["bf16_tflops"], tf_sustained=d.get("bf16_tflops_sustained", d["bf16_tflops"]),
                    source="measured")
    return dict(hbm=6650.0, tf_burst=1590.0, tf_sustained=1400.0, source="fallback")


class ClockSampler(threading.Thread):
    """nvidia-smi clocks / throttle reasons during the timed region (B200_PROFILING.md recipe)."""

    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index: int):
        super().__init__(daemon=True)
        self.index, self.samples, self._stop_evt = index, [], threading.Event()

    def run(self):
        while not self._stop_evt.is_set():
            try:
                out = subprocess.run(["nvidia-smi", f"--id={self.index}", f"--query-gpu={self.Q}",
                                      "--format=csv,noheader,nounits"], capture_output=True, text=True, timeout=5).stdout
                parts = [p.strip() for p in out.strip().split(",")]
                if len(parts) >= 7:
                    self.samples.append(parts)
            except Exception:
                pass
            self._stop_evt.wait(0.2)

    def stop(self):
        self._stop_evt.set()
        self.join(timeout=3)
        sm = sorted(float(s[0]) for s in self.samples if s[0].replace(".", "").isdigit())
        reasons = set()
        for s in self.samples:
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), s[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": sm[len(sm) // 2] if sm else None,
                "sm_max_mhz": float(self.samples[0][1]) if self.samples else None,
                "reasons": sorted(reasons), "samples": len(self.samples)}


def chunk_parallelism(world):
    return f"dp{world}: one independent T-frame chunk per rank, no data-path collective"


def workload_config(T, h, w, parallelism):
    """`config` of both arms (identical workload string: the driver compares them)."""
    return {"workload": f"Seva 1.3B random-init, T={T} @{h * 8}x{w * 8} (latent {h}x{w}), Euler-EDM step with "
                        f"MultiviewCFG (CFG batch {2 * T}), configs[1]",
            "parallelism": parallelism,
            "l2": "activations per step (>2 GB) exceed the 126 MB L2; no explicit flush"}


# ----------------------------------------------------------------------------------------------------
def build_model(device):
    from seva_b200.model import Seva, SevaParams, SGMWrapper

    params = SevaParams()
    with torch.device("meta"):
        m = Seva(params)
    g = torch.Generator(device=device).manual_seed(0)
    sd = {}
    for k, v in m.state_dict().items():  # random-init, default PyTorch scale, generated on the device
        if v.ndim >= 2:
            t = (torch.rand(v.shape, device=device, generator=g) * 2 - 1) / math.sqrt(math.prod(v.shape[1:]))
        elif k.endswith("weight"):
            t = 1.0 + 0.1 * torch.randn(v.shape, device=device, generator=g)
        else:
            t = 0.05 * torch.randn(v.shape, device=device, generator=g)
        sd[k] = t
    m.load_state_dict(sd, assign=True)
    return SGMWrapper(m.eval()), params


def make_host_inputs(T, h, w, seed):
    """Synthetic chunk in pinned host memory."""
    from oracle import synth  # input generator only (shared with the tests); no oracle compute on this path
    from seva_b200 import geometry as G

    inp = synth.make_inputs(T, 1, h, w, lambda s, e, K, hh, ww: G.get_plucker_coordinates(s, e, K, target_size=(hh, ww)),
                            seed=seed, num_steps=1)
    pin = lambda t: t.contiguous().pin_memory()
    return dict(x0=pin(inp["x0"]), c={k: pin(v) for k, v in inp["c"].items()}, uc={k: pin(v) for k, v in inp["uc"].items()},
                c2w=pin(inp["c2w"]), K=pin(inp["K"]), mask=pin(inp["mask"]))


def strong_scaling_extra(args, model, dev, world, rank):
    """Extra key of the N > 1 line (the headline `value` stays chunk data-parallel, the metric's workload): ONE T-frame
    chunk over all N GPUs -- CFG halves on rank pairs x frame sequence-parallel over N/2 (BASELINE configs[3]) -- so
    that the driver's scaling file also records the strong-scaling curve and the K/V-exchange share.  Collective: every
    rank must call it.  Device time (CUDA events), max over ranks."""
    import torch.distributed as dist

    from seva_b200 import dist as D
    from seva_b200 import sampling as S
    from seva_b200.engine import _FrameGather, _RingAttention, _RingStart, _SpaceExchange

    T, h, w = args.T, args.size, args.size
    topo = D.Topology(world, rank, cfg=2, sp=world // 2)
    groups = D.build_groups(topo)
    if groups["frames"] is not None:
        model.module.set_frame_group(groups["frames"])
    guider = D.CFGSplit(S.MultiviewCFG(1.2), groups["cfg"], half=topo.half)
    torch.manual_seed(1234)  # replicated sampler state: identical noise on every rank
    host = make_host_inputs(T, h, w, seed=1)
    disc = S.DDPMDiscretization()
    denoiser = S.DiscreteDenoiser(disc, num_idx=1000, device=dev)
    sampler = S.EulerEDMSampler(disc, guider, num_steps=50, s_churn=0.0, s_tmin=0.0, s_tmax=999.0, s_noise=1.0,
                                verbose=False, device=dev)
    sigmas = disc(50, device=dev)
    nb = lambda t: t.to(dev)
    x = nb(host["x0"]) * 84.9222
    c, uc = {k: nb(v) for k, v in host["c"].items()}, {k: nb(v) for k, v in host["uc"].items()}
    kw = dict(c2w=nb(host["c2w"]), K=nb(host["K"]), input_frame_mask=nb(host["mask"]))
    den_fn = lambda i, s_, cc: denoiser(model, i, s_, cc, num_frames=T)
    ones = x.new_ones([T])
    step = lambda i: sampler.sampler_step(ones * sigmas[i % 50], ones * sigmas[i % 50 + 1], den_fn, x, 2.0, c, uc, 0.0, **kw)
    steps = max(3, min(args.steps, 10))
    with torch.inference_mode(), torch.autocast("cuda"):
        for i in range(3):
            step(i)
        dist.barrier()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for i in range(steps):
            out = step(i)
        e1.record()
        dist.barrier()
        torch.cuda.synchronize()
        t = torch.tensor([e0.elapsed_time(e1)], device=dev, dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        # exchange share: the same steps once more with an event pair around every plan entry (rank 0's view)
        eng = next(iter(model.module._engines.values()))
        timing: list = []
        eng.timing_sink = timing
        for i in range(2):
            step(i)
        torch.cuda.synchronize()
        eng.timing_sink = None
    xch = sum(a.elapsed_time(b) for idx, a, b in timing
              if isinstance(eng.calls[idx][0], (_FrameGather, _SpaceExchange, _RingStart))) / 2
    ring = sum(a.elapsed_time(b) for idx, a, b in timing if isinstance(eng.calls[idx][0], _RingAttention)) / 2
    total = sum(a.elapsed_time(b) for idx, a, b in timing) / 2
    assert torch.isfinite(out).all()
    model.module.set_frame_group(None)
    ms = t.item() / steps
    return {"value": round(T / (ms / 1e3), 3), "unit": "frame-steps/s", "ms_per_step": round(ms, 3), "steps": steps,
            "scaling": "strong", "parallelism": f"one chunk over {world} GPUs: cfg2 x frames{world // 2}",
            "blocking_exchange_ms_per_step": round(xch, 3), "attn3d_with_overlapped_exchange_ms_per_step": round(ring, 3),
            "plan_ms_per_step": round(total, 3), "ring_mode": os.environ.get("SEVA_RING", "ring")}


def run_ours(args):
    from seva_b200 import sampling as S

    rank = int(os.environ.get("RANK", 0))
    world = int(os.environ.get("WORLD_SIZE", 1))
    local = int(os.environ.get("LOCAL_RANK", 0))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        import torch.distributed as dist

        dist.init_process_group("nccl", device_id=dev)
    T, h, w = args.T, args.size, args.size
    model, _ = build_model(dev)
    # --parallel chunks (default): every rank samples its own chunk (weak scaling, no data-path collective)
    # --parallel cfg:    ONE chunk; uncond / cond halves on rank pairs, remaining factor world/2 splits the frames
    # --parallel frames: ONE chunk; all ranks split the frames (K/V exchanged at 3D / time-mix attention)
    guider = S.MultiviewCFG(1.2)
    chunks_world, par_desc = world, chunk_parallelism(world)
    if world > 1 and args.parallel != "chunks":
        from seva_b200 import dist as D

        cfg = 2 if args.parallel == "cfg" else 1
        topo = D.Topology(world, rank, cfg=cfg, sp=world // cfg)
        groups = D.build_groups(topo)
        if groups["frames"] is not None:
            model.module.set_frame_group(groups["frames"])
        if cfg == 2:
            guider = D.CFGSplit(guider, groups["cfg"], half=topo.half)
        chunks_world = 1
        par_desc = f"one chunk over {world} GPUs: cfg{cfg} x frames{world // cfg}"
        torch.manual_seed(1234)  # replicated sampler state: identical noise on every rank
    host = make_host_inputs(T, h, w, seed=1 + (rank if chunks_world > 1 else 0))
    disc = S.DDPMDiscretization()
    denoiser = S.DiscreteDenoiser(disc, num_idx=1000, device=dev)
    sampler = S.EulerEDMSampler(disc, guider, num_steps=50, s_churn=0.0, s_tmin=0.0, s_tmax=999.0,
                                s_noise=1.0, verbose=False, device=dev)
    sigmas = disc(50, device=dev)
    den_fn = lambda i, s, c: denoiser(model, i, s, c, num_frames=T)

    def upload():
        nb = lambda t: t.to(dev, non_blocking=True)
        return (nb(host["x0"]), {k: nb(v) for k, v in host["c"].items()}, {k: nb(v) for k, v in host["uc"].items()},
                nb(host["c2w"]), nb(host["K"]), nb(host["mask"]))

    h2d = (host["x0"].numel() * 4 + sum(v.numel() * v.element_size() for d in (host["c"], host["uc"]) for v in d.values())
           + host["c2w"].numel() * 4 + host["K"].numel() * 4 + host["mask"].numel())
    d2h = host["x0"].numel() * 4
    x_host = torch.empty_like(host["x0"]).pin_memory()

    def step(i, x, c, uc, c2w, K, mask):
        ones = x.new_ones([T])
        k = i % 50
        return sampler.sampler_step(ones * sigmas[k], ones * sigmas[k + 1], den_fn, x, 2.0, c, uc, 0.0, c2w=c2w, K=K,
                                    input_frame_mask=mask)

    def barrier():
        if world > 1:
            import torch.distributed as dist

            dist.barrier()
        torch.cuda.synchronize()

    def maxed(ms):
        if world > 1:
            import torch.distributed as dist

            t = torch.tensor([ms], device=dev, dtype=torch.float64)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            return t.item()
        return ms

    with torch.inference_mode(), torch.autocast("cuda"):
        x, c, uc, c2w, K, mask = upload()
        x = x * 84.9222
        for i in range(args.warmup):
            xw = step(i, x, c, uc, c2w, K, mask)
        barrier()
        eng = next(iter(model.module._engines.values()))
        # ---- device-resident timing (value): K steps, one CUDA-event pair, nothing else on the stream ---------
        clocks = ClockSampler(local)
        clocks.start()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        barrier()
        e0.record()
        xs = x
        for i in range(args.steps):
            xs = step(i, xs, c, uc, c2w, K, mask)
        e1.record()
        barrier()
        ms_dev = maxed(e0.elapsed_time(e1))
        # ---- the same K steps again with a CUDA-event pair around every launch (launching stream): per-kernel
        # durations for the roofline.  Kept out of `value`: ~1200 extra event records per step cost 1-2 %.
        timing: list = []
        eng.timing_sink = timing
        barrier()
        e0.record()
        xs = x
        for i in range(args.steps):
            xs = step(i, xs, c, uc, c2w, K, mask)
        e1.record()
        barrier()
        clk = clocks.stop()
        ms_inst = maxed(e0.elapsed_time(e1))
        eng.timing_sink = None
        # ---- end-to-end timing: host buffers in, host result out, every step ---------------------
        # Every step's inputs come from pinned host memory and its result goes back to the host, all inside the timed
        # region.  Two preallocated device input sets: the upload of step i + 1 runs on a copy stream while step i
        # computes (what a serving loop does); the compute stream waits for its own step's upload, the copy stream for the
        # step that last read the set it overwrites.  The conditioning changes identity every step, so the guider's
        # prepared-input cache and the engine's step-invariant hoisting are OFF on this path (every step pays them).
        main = torch.cuda.current_stream()
        copy_stream = torch.cuda.Stream()
        flat = lambda t6: [t6[0], *t6[1].values(), *t6[2].values(), *t6[3:]]
        host_flat = flat((host["x0"], host["c"], host["uc"], host["c2w"], host["K"], host["mask"]))
        sets = [upload() for _ in range(2)]
        torch.cuda.synchronize()
        uploaded = [torch.cuda.Event() for _ in range(2)]
        consumed = [torch.cuda.Event() for _ in range(2)]

        def upload_into(k, first):
            with torch.cuda.stream(copy_stream):
                if not first:
                    copy_stream.wait_event(consumed[k])
                for dst, src in zip(flat(sets[k]), host_flat):
                    dst.copy_(src, non_blocking=True)
                uploaded[k].record(copy_stream)

        def e2e_steps(n):
            upload_into(0, True)
            for i in range(n):
                k = i % 2
                main.wait_event(uploaded[k])
                if i + 1 < n:
                    upload_into(1 - k, i == 0)
                xh, c_, uc_, c2w_, K_, mask_ = sets[k]
                xn = step(i, xh * 84.9222, c_, uc_, c2w_, K_, mask_)
                consumed[k].record(main)
                x_host.copy_(xn, non_blocking=True)

        e2e_steps(2)  # untimed: this path's own launch variant (no hoisting) is captured into its CUDA graph here
        barrier()
        e0.record()
        e2e_steps(args.steps)
        e1.record()
        barrier()
        ms_e2e = maxed(e0.elapsed_time(e1))
        assert torch.isfinite(x_host).all()

    # ---- per-kernel aggregation from the live events -----------------------------------------------
    agg: dict = {}
    for idx, s0, s1 in timing:
        name, tag, fl, nb = eng.meta[idx]
        a = agg.setdefault(name, dict(ms=0.0, flops=0.0, bytes=0.0, n=0))
        a["ms"] += s0.elapsed_time(s1)
        a["flops"] += fl
        a["bytes"] += nb
        a["n"] += 1
    if args.dump_launches and rank == 0:
        rows = [dict(i=idx, name=eng.meta[idx][0], tag=eng.meta[idx][1], flops=eng.meta[idx][2], bytes=eng.meta[idx][3],
                     ms=s0.elapsed_time(s1)) for idx, s0, s1 in timing]
        os.makedirs(os.path.dirname(os.path.abspath(args.dump_launches)), exist_ok=True)
        with open(args.dump_launches, "w") as f:
            json.dump(rows, f)
    peaks = measured_peaks()
    traffic = {}
    tpath = os.path.join(ROOT, "profiles", "ncu_traffic.json")  # DRAM bytes per launch from the committed ncu pass
    if os.path.exists(tpath) and T == 21 and h == 72 and w == 72:
        with open(tpath) as f:
            traffic = json.load(f).get("kernels", {})
    plan_ms = sum(a["ms"] for a in agg.values())
    top = max(agg.items(), key=lambda kv: kv[1]["ms"])
    kernels = {k: dict(share=round(a["ms"] / plan_ms, 4), ms_per_step=round(a["ms"] / args.steps, 3),
                       launches_per_step=a["n"] // args.steps,
                       tflops=round(a["flops"] / a["ms"] / 1e9, 1) if a["flops"] and a["ms"] else None,
                       gbs=round(a["bytes"] / a["ms"] / 1e6, 1) if a["bytes"] and a["ms"] and not a["flops"] else None)
               for k, a in sorted(agg.items(), key=lambda kv: -kv[1]["ms"])}
    tname, ta = top
    if ta["flops"]:
        ach = ta["flops"] / ta["ms"] / 1e9
        roof = dict(kernel=tname, bound="tensor", achieved=round(ach, 1), peak=peaks["tf_sustained"], unit="TFLOP/s",
                    frac=round(ach / peaks["tf_sustained"], 4), traffic=traffic.get(tname),
                    traffic_source="profiles/ncu_traffic.json (ncu dram__bytes_read+write per launch, mean)" if tname in traffic else None,
                    peak_source=peaks["source"] + " sustained bf16 (kernel timed inside a long step)",
                    launches=ta["n"], avg_launch_ms=round(ta["ms"] / ta["n"], 4),
                    algorithmic_flops_per_launch=ta["flops"] / ta["n"])
    else:
        ach = ta["bytes"] / ta["ms"] / 1e6
        roof = dict(kernel=tname, bound="hbm", achieved=round(ach, 1), peak=peaks["hbm"], unit="GB/s",
                    frac=round(ach / peaks["hbm"], 4), traffic=traffic.get(tname), peak_source=peaks["source"],
                    launches=ta["n"], avg_launch_ms=round(ta["ms"] / ta["n"], 4))

    n_launch_step = eng.n_launch_step
    strong = None
    if world > 1 and world % 2 == 0 and args.parallel == "chunks" and not args.no_strong:
        del eng
        model.module.invalidate()
        torch.cuda.empty_cache()
        try:
            strong = strong_scaling_extra(args, model, dev, world, rank)
        except Exception as e:  # never take the headline line down
            strong = {"unavailable": f"{type(e).__name__}: {e}"[:300]}
    if rank != 0:
        return
    steps = args.steps
    value = chunks_world * T * steps / (ms_dev / 1e3)
    e2e_v = chunks_world * T * steps / (ms_e2e / 1e3)
    tf_step = REF_TFLOP_PER_STEP.get(T)
    lib = None
    if world == 1 and not args.no_gpu_baseline:
        model.module.invalidate()  # return our workspaces before the reference modules allocate theirs
        del eng
        torch.cuda.empty_cache()
        try:
            lib = gpu_library_baseline(T, h, w, dev)
        except Exception as e:  # a baseline leg must never take the product line down
            lib = {"unavailable": f"{type(e).__name__}: {e}"[:300]}
        torch.cuda.empty_cache()
    cpu = cpu_baseline(args) if world == 1 and not args.no_cpu else None
    line = {
        "metric": "frame-steps/s", "value": round(value, 3), "unit": "frame-steps/s", "n_gpus": world, "steps": steps,
        "warmup": args.warmup, "ms_per_step": round(ms_dev / steps, 3), "higher_is_better": True,
        "scaling": "weak" if chunks_world == world else "strong", "vs_baseline": None, "dtype": "bf16", "data": "synthetic",
        "config": workload_config(T, h, w, par_desc),
        "e2e": {"value": round(e2e_v, 3), "unit": "frame-steps/s", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                "ms_per_step": round(ms_e2e / steps, 3)},
        "gpu_launches": n_launch_step * steps + 4 * steps,
        "clocks": clk, "roofline": roof, "ms_per_step_with_launch_events": round(ms_inst / steps, 3),
        "step_tflops": round(chunks_world * tf_step * steps / (ms_dev / 1e3), 1) if tf_step else None,
        "step_frac_of_peak": round(chunks_world * tf_step * steps / (ms_dev / 1e3) / peaks["tf_sustained"] / world, 4)
        if tf_step else None,
        "kernels": kernels,
    }
    if cpu is not None:
        line["cpu_baseline"] = cpu
    if lib is not None:
        line["gpu_library_baseline"] = lib
    if strong is not None:
        line["strong_cfg_frames"] = strong
    print(json.dumps(line))


# ----------------------------------------------------------------------------------------------------
def arc_cameras(n, seed):
    """[n,4,4] c2w along a jittered arc (synthetic trajectory for the scene workload)."""
    g = torch.Generator().manual_seed(seed)
    t = torch.linspace(0, 1, n)
    yaw = 1.5 * t + 0.05 * torch.randn(n, generator=g)
    c2w = torch.eye(4).repeat(n, 1, 1)
    c2w[:, 0, 0], c2w[:, 0, 2], c2w[:, 2, 0], c2w[:, 2, 2] = torch.cos(yaw), torch.sin(yaw), -torch.sin(yaw), torch.cos(yaw)
    c2w[:, :3, 3] = torch.stack([2 * torch.sin(2.5 * t), 0.3 * torch.cos(4 * t), -2 * torch.cos(2.5 * t)], 1) \
        + 0.05 * torch.randn(n, 3, generator=g)
    return c2w


def run_scene_bench(args):
    """BASELINE.json configs[2]: two-pass chunked sampling, 1 input -> 80 target views, T=21 windows (1 anchor window +
    7 target windows = 8 runs x 21 frames x 50 steps = 8400 frame-steps), scheduled over the N GPUs by
    seva_b200.scene: the anchor window model-parallel over all ranks (CFG halves x frame shards), the 7 independent
    target windows dealt round-robin.  VAE / CLIP are outside the path: identity codec, fixed token projection."""
    import numpy as np

    from seva_b200 import dist as D
    from seva_b200 import planner as P
    from seva_b200 import sampling as S
    from seva_b200 import scene as SC

    rank, world = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
    local = int(os.environ.get("LOCAL_RANK", 0))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    where = SC.Rank()
    if world > 1:
        import torch.distributed as dist

        dist.init_process_group("nccl", device_id=dev)
        topo = D.Topology(world, rank, cfg=2, sp=world // 2)
        groups = D.build_groups(topo)
        where = SC.Rank(rank=rank, world=world, chain_frames=groups["frames"],
                        chain_guider=lambda g: D.CFGSplit(g, groups["cfg"], half=topo.half))
    model, _ = build_model(dev)
    T, hh, n_targets = args.T, args.size, 80
    f = 0.5 / math.tan(0.5 * 0.9424777960769379)
    Kn = lambda n: torch.tensor([[f, 0, 0.5], [0, f, 0.5], [0, 0, 1.0]]).repeat(n, 1, 1)
    traj, inp = arc_cameras(n_targets, 3), arc_cameras(1, 4)
    opts = {"chunk_strategy": "interp", "chunk_strategy_first_pass": "gt-nearest", "cfg": [3.0, 2.0]}
    version = {"T": [T, T], "options": opts}
    prior = P.infer_prior_inds(traj, P.infer_prior_stats([T, T], 1, n_targets, version), [], opts)
    sc = SC.Scene(torch.randn(1, 4, hh, hh, generator=torch.Generator().manual_seed(5)), inp, Kn(1), traj, Kn(n_targets),
                  [0], list(range(1, n_targets + 1)), traj[prior], Kn(len(prior)), [int(i) + 1 for i in prior])
    plan = SC.plan_scene(sc, version["T"], opts, task="img2trajvid", use_traj_prior=True)
    n_runs = sum(len(p.windows) for p in plan.passes)
    disc = S.DDPMDiscretization()
    denoiser = S.DiscreteDenoiser(disc, num_idx=1000, device=dev)

    def samplers(steps):
        def make(T_, wrap):
            g = S.MultiviewCFG(1.2)
            return S.EulerEDMSampler(disc, wrap(g) if wrap else g, num_steps=steps, s_churn=0.0, s_tmin=0.0,
                                     s_tmax=999.0, s_noise=1.0, verbose=False, device=dev)
        return make

    def sync():
        if world > 1:
            import torch.distributed as dist

            dist.barrier()
        torch.cuda.synchronize()

    torch.manual_seed(1234)
    SC.run_scene(sc, plan, model, denoiser, samplers(2), where=where, generator=torch.Generator().manual_seed(7),
                 device=dev)  # warm-up: builds every plan (sharded and unsharded), 2 steps per run
    clocks = ClockSampler(local)
    clocks.start()
    sync()
    t0 = time.time()
    out = SC.run_scene(sc, plan, model, denoiser, samplers(50), where=where,
                       generator=torch.Generator().manual_seed(7), device=dev)
    sync()
    secs = time.time() - t0
    clk = clocks.stop()
    if world > 1:
        import torch.distributed as dist

        t = torch.tensor([secs], device=dev, dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        secs = t.item()
    assert out.shape[0] == n_targets and torch.isfinite(out).all()
    if rank == 0:
        fs = n_runs * T * 50
        print(json.dumps({
            "metric": "frame-steps/s", "value": round(fs / secs, 3), "unit": "frame-steps/s", "n_gpus": world,
            "steps": 50 * n_runs, "warmup": 2 * n_runs, "seconds": round(secs, 3), "higher_is_better": True,
            "scaling": "strong", "vs_baseline": None, "dtype": "bf16", "data": "synthetic",
            "config": {"workload": f"two-pass scene, 1 input -> {n_targets} targets, T={T} @{hh * 8}x{hh * 8}: "
                                   f"{len(plan.passes[0].windows)} anchor + {len(plan.passes[1].windows)} target windows x "
                                   f"50 Euler-EDM steps with MultiviewCFG = {fs} frame-steps, configs[2]",
                       "parallelism": "single GPU" if world == 1 else
                       f"anchor window over cfg2 x frames{world // 2}; target windows round-robin over {world} GPUs",
                       "timing": "host wall clock around the whole scene (barrier + synchronize both sides), max over ranks"},
            "clocks": clk, "latent_checksum": round(float(out.double().abs().mean()), 6)}))


def _reference_rig(device, T, h, w):
    """The REFERENCE's own objects for the path (oracle/_ref staged by oracle/make_ref.py, or /root/reference):
    SGMWrapper(Seva(SevaParams())) with default-init weights (seed 0), DiscreteDenoiser, EulerEDMSampler with
    MultiviewCFG(1.2), and the synthetic chunk of SURVEY.md 8d.  Returns None when neither copy exists."""
    from oracle import refimport, synth

    if not (refimport.available() or refimport.staged_available()):
        return None
    ref = refimport.import_staged()
    M, S, G = ref.model, ref.sampling, ref.geometry
    torch.manual_seed(0)
    net = M.SGMWrapper(M.Seva(M.SevaParams()).eval()).to(device)
    disc = S.DDPMDiscretization()
    denoiser = S.DiscreteDenoiser(disc, num_idx=1000, device=device)
    sampler = S.EulerEDMSampler(disc, S.MultiviewCFG(1.2), num_steps=50, s_churn=0.0, s_tmin=0.0, s_tmax=999.0,
                                s_noise=1.0, verbose=False, device=device)
    plk = lambda s_, e_, K_, hh, ww: G.get_plucker_coordinates(extrinsics_src=s_, extrinsics=e_, intrinsics=K_,
                                                               target_size=(hh, ww))
    inp = synth.make_inputs(T, 1, h, w, plk, seed=1, num_steps=1)
    mv = lambda d: {k: v.to(device) for k, v in d.items()}
    sig = disc(50, device=device)
    ones = torch.ones(T, device=device)
    x = inp["x0"].to(device) * 84.9222
    c, uc = mv(inp["c"]), mv(inp["uc"])
    kw = dict(c2w=inp["c2w"].to(device), K=inp["K"].to(device), input_frame_mask=inp["mask"].to(device))
    den = lambda i_, s_, c_: denoiser(net, i_, s_, c_, num_frames=T)

    def step(i):
        k = i % 50
        return sampler.sampler_step(ones * sig[k], ones * sig[k + 1], den, x, 2.0, c, uc, 0.0, **kw)

    return step, ref.root


def _port_rig(T, h, w):
    """Fallback when the reference copy is absent: the oracle PORT (plain PyTorch fp32 restatement)."""
    from oracle import seva_oracle as O
    from oracle import synth
    from seva_b200.model import Seva, SevaParams

    P = SevaParams()
    with torch.device("meta"):
        m = Seva(P)
    shapes = {k: tuple(v.shape) for k, v in m.state_dict().items()}
    g = torch.Generator().manual_seed(0)
    sd = {k: ((torch.rand(s, generator=g) * 2 - 1) / math.sqrt(math.prod(s[1:])) if len(s) >= 2 else
              (torch.ones(s) if k.endswith("weight") else torch.zeros(s))) for k, s in shapes.items()}
    inp = synth.make_inputs(T, 1, h, w, O.plucker, seed=1, num_steps=1)
    table = O.denoiser_table()
    fs = O.multiview_scale(2.0, 1.2, inp["c2w"], inp["K"], inp["mask"])
    sig = torch.cat([O.ddpm_sigmas(50), torch.zeros(1)])
    ones = torch.ones(T)
    net = lambda x, idx, c: O.sgm_forward(sd, P, x, idx, c, num_frames=T)

    def step(i):
        k = i % 50
        return O.sampler_step(net, table, inp["x0"] * 84.9222, ones * sig[k], ones * sig[k + 1], fs, inp["c"], inp["uc"],
                              inp["eps"][0])[0]

    return step


def cpu_reference_run(T, h, w, steps, warmup, budget_s):
    """Times the reference's CPU implementation of the path on all host cores, fp32: ``EulerEDMSampler.sampler_step``
    of the staged reference at the FULL workload (T frames, CFG batch 2T).  One step costs minutes at T=21, so the
    run is bounded: at most ``steps`` timed steps, stopping once ``budget_s`` is spent (always at least one); warm-up
    steps only while they fit the same budget.  Returns (seconds per step, steps timed, warm-ups done, kind, cores)."""
    threads = os.cpu_count() or 1
    torch.set_num_threads(threads)
    rig = _reference_rig(torch.device("cpu"), T, h, w)
    kind = "reference"
    if rig is None:
        step, kind = _port_rig(T, h, w), "port"
    else:
        step = rig[0]
    times, n_warm, out = [], 0, None
    t_start = time.time()
    with torch.inference_mode():
        est = None
        for i in range(warmup):
            if est is not None and (time.time() - t_start) + 2 * est > budget_s:
                break
            t0 = time.time()
            out = step(i)
            est = time.time() - t0
            if (time.time() - t_start) + est > budget_s:  # a warm-up that would eat the only timed step: count it as timed
                times.append(est)
                break
            n_warm += 1
        for i in range(steps):
            if times and (time.time() - t_start) + times[-1] > budget_s:
                break
            t0 = time.time()
            out = step(n_warm + i)
            times.append(time.time() - t0)
    assert out is not None and torch.isfinite(out).all()
    return sum(times) / len(times), len(times), n_warm, kind, threads


def cpu_baseline(args):
    """`cpu_baseline` of the product line: a bounded sample (ONE step) of the same workload on the host cores."""
    secs, n, _, kind, threads = cpu_reference_run(args.T, args.size, args.size, 1, 0, args.cpu_budget)
    what = ("the reference's own EulerEDMSampler.sampler_step (oracle/_ref copy of seva/{model,sampling,geometry}.py + "
            "modules)" if kind == "reference" else "the oracle port (reference copy absent)")
    return {"value": round(args.T / secs, 5), "unit": "frame-steps/s", "cores": threads, "kind": kind,
            "sample": f"{n} Euler-EDM step with MultiviewCFG of {what}, PyTorch fp32, {threads} threads, at the full "
                      f"workload T={args.T} @{args.size * 8}x{args.size * 8} (CFG batch {2 * args.T}); {secs:.1f} s/step"}


def gpu_library_baseline(T, h, w, device, steps=3):
    """the reference modules themselves on the SAME B200 in eager PyTorch under bf16
    autocast (library kernels: cuBLAS / cuDNN / the flash SDPA the reference forces, transformer.py:70) - the honest
    GPU baseline next to our kernels.  One warm-up step, then ``steps`` steps between CUDA events."""
    rig = _reference_rig(device, T, h, w)
    if rig is None:
        return {"unavailable": "reference copy (oracle/_ref) not staged"}
    step = rig[0]
    with torch.inference_mode(), torch.autocast("cuda", dtype=torch.bfloat16):
        step(0)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for i in range(steps):
            out = step(1 + i)
        e1.record()
        torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / steps
    assert torch.isfinite(out).all()
    return {"value": round(T / (ms / 1e3), 3), "unit": "frame-steps/s", "ms_per_step": round(ms, 2), "steps": steps,
            "what": "UNMODIFIED reference modules (oracle/_ref: seva.model / seva.sampling), eager PyTorch, "
                    "torch.autocast(bf16), forced FLASH SDPA, same GPU, same workload, inputs resident in HBM"}


def run_reference(args):
    rank = int(os.environ.get("RANK", 0))
    if rank != 0:
        return
    secs, n, n_warm, kind, threads = cpu_reference_run(args.T, args.size, args.size, max(1, args.steps), args.warmup,
                                                       args.cpu_budget)
    v = round(args.T / secs, 5)
    world = int(os.environ.get("WORLD_SIZE", 1))
    h = w = args.size
    what = ("reference's own code (oracle/_ref: seva.sampling.EulerEDMSampler.sampler_step -> DiscreteDenoiser -> "
            "SGMWrapper(Seva) -> MultiviewCFG)" if kind == "reference" else "oracle port (reference copy absent)")
    print(json.dumps({
        "impl": "reference", "metric": "frame-steps/s", "value": v, "unit": "frame-steps/s", "n_gpus": world,
        "steps": n, "warmup": n_warm, "ms_per_step": round(secs * 1e3, 1), "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": workload_config(args.T, h, w, chunk_parallelism(world)),  # identical to the product arm's at this N
        "cpu_baseline": {"value": v, "unit": "frame-steps/s", "cores": threads, "kind": kind,
                         "sample": f"{what}, PyTorch fp32 on {threads} host threads at the full workload T={args.T} "
                                   f"(CFG batch {2 * args.T}); {n} timed step(s) of {secs:.1f} s after {n_warm} warm-up "
                                   f"(bounded to ~{args.cpu_budget:.0f} s: requested --steps {args.steps} --warmup {args.warmup})"},
        "e2e": {"value": v, "unit": "frame-steps/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }))


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--T", type=int, default=21)
    ap.add_argument("--size", type=int, default=72)
    ap.add_argument("--parallel", default="chunks", choices=["chunks", "cfg", "frames"],
                    help="how N > 1 GPUs are used (see run_ours); the driver's scaling run uses the default")
    ap.add_argument("--workload", default="step", choices=["step", "scene"],
                    help="step: one Euler-EDM step of configs[1] (the driver's contract); scene: configs[2] end to end")
    ap.add_argument("--no-cpu", action="store_true", help="skip the CPU baseline leg")
    ap.add_argument("--no-strong", action="store_true", help="N > 1: skip the extra strong-scaling (cfg2 x frames) leg")
    ap.add_argument("--no-gpu-baseline", action="store_true", help="skip the eager-PyTorch reference-on-GPU leg")
    ap.add_argument("--cpu-budget", type=float, default=float(os.environ.get("SEVA_CPU_BUDGET_S", 300)),
                    help="seconds the CPU reference leg may spend (it always times at least one full step)")
    ap.add_argument("--dump-launches", default=None, help="write per-launch CUDA-event timings (JSON) here")
    a = ap.parse_args()
    if a.impl == "reference":
        run_reference(a)
    else:
        if not torch.cuda.is_available():
            raise SystemExit("bench.py --impl ours needs a B200: the product path has no CPU fallback")
        run_scene_bench(a) if a.workload == "scene" else run_ours(a)
