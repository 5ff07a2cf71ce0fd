"""Host logic of the multi-GPU partitioning (seva_b200/dist.py) on CPU: integer bookkeeping bit-exact, and the
collective plumbing over a world_size-2 gloo group (the same code runs over NCCL on the GPUs)."""
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from seva_b200 import dist as D
from seva_b200 import sampling as S


def test_partition_frames():
    assert D.partition_frames(21, 1) == [(0, 21)]
    assert D.partition_frames(21, 2) == [(0, 11), (11, 21)]
    assert D.partition_frames(21, 4) == [(0, 6), (6, 11), (11, 16), (16, 21)]
    assert [b - a for a, b in D.partition_frames(21, 8)] == [3, 3, 3, 3, 3, 2, 2, 2]
    assert D.partition_frames(64, 8) == [(8 * r, 8 * r + 8) for r in range(8)]
    assert D.partition_frames(2, 4) == [(0, 1), (1, 2), (2, 2), (2, 2)]
    for T in (1, 5, 21, 64):
        for P in (1, 2, 3, 4, 8):
            parts = D.partition_frames(T, P)
            assert parts[0][0] == 0 and parts[-1][1] == T
            assert all(parts[i][1] == parts[i + 1][0] for i in range(P - 1))
    with pytest.raises(ValueError):
        D.partition_frames(0, 2)


def test_schedule_chunks():
    # 1 first-pass + 7 second-pass chunks; the 7 independent ones over 8 / 4 / 2 GPUs
    assert D.schedule_chunks(7, 8) == [[0], [1], [2], [3], [4], [5], [6], []]
    assert D.schedule_chunks(7, 4) == [[0, 4], [1, 5], [2, 6], [3]]
    assert D.schedule_chunks(7, 2) == [[0, 2, 4, 6], [1, 3, 5]]
    assert D.schedule_chunks(0, 3) == [[], [], []]
    flat = sorted(k for r in D.schedule_chunks(11, 3) for k in r)
    assert flat == list(range(11))


def test_topology():
    t = D.Topology(world=8, rank=5, cfg=2, sp=2)
    assert (t.chunk_rank, t.half, t.frame_rank, t.chunk_world) == (1, 0, 1, 2)
    assert t.frame_group_ranks() == [4, 5] and t.cfg_group_ranks() == [5, 7] and t.model_group_ranks() == [4, 5, 6, 7]
    seen = set()
    for r in range(8):
        tt = D.Topology(8, r, 2, 2)
        seen.add((tt.chunk_rank, tt.half, tt.frame_rank))
    assert len(seen) == 8
    with pytest.raises(ValueError):
        D.Topology(world=6, rank=0, cfg=2, sp=2)
    with pytest.raises(ValueError):
        D.Topology(world=4, rank=0, cfg=3, sp=1)


# ---- world_size-2 gloo workers --------------------------------------------------------------------
def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _fake_denoiser(x, sigma, cond):
    """Deterministic stand-in for the network: depends on every conditioning tensor and on the batch position, so a
    wrong half / wrong order cannot cancel out."""
    cond.pop("replace", None)
    ca = cond["crossattn"].mean(dim=(1, 2)).view(-1, 1, 1, 1)
    cc = cond["concat"][:, :4]
    return x * 0.5 + 0.1 * ca + 0.05 * cc + 0.01 * sigma.view(-1, 1, 1, 1)


def _sampling_inputs(T=5, h=8, w=8):
    g = torch.Generator().manual_seed(3)
    r = lambda *s: torch.randn(*s, generator=g)
    c = {"crossattn": r(T, 1, 16), "replace": r(T, 5, h, w), "concat": r(T, 7, h, w), "dense_vector": r(T, 6, h, w)}
    uc = {"crossattn": torch.zeros(T, 1, 16), "replace": torch.zeros(T, 5, h, w), "concat": r(T, 7, h, w),
          "dense_vector": c["dense_vector"].clone()}
    c2w = torch.eye(4).repeat(T, 1, 1)
    c2w[:, 0, 3] = torch.arange(T).float()
    K = torch.eye(3).repeat(T, 1, 1)
    mask = torch.zeros(T, dtype=torch.bool)
    mask[0] = True
    return r(T, 4, h, w), c, uc, c2w, K, mask


def _run_sampler(guider, x0, c, uc, c2w, K, mask, steps=4):
    disc = S.DDPMDiscretization()
    sampler = S.EulerEDMSampler(disc, guider, num_steps=steps, s_churn=0.0, s_tmin=0.0, s_tmax=999.0, s_noise=1.0,
                                verbose=False, device="cpu")
    torch.manual_seed(11)  # identical noise on every rank
    return sampler(_fake_denoiser, x0.clone(), scale=2.0, cond={k: v.clone() for k, v in c.items()},
                   uc={k: v.clone() for k, v in uc.items()}, verbose=False, c2w=c2w, K=K, input_frame_mask=mask)


class _OracleFrameGather:
    """sp hook of oracle.seva_forward over a real process group: K/V of the other ranks' frames via allgather_rows."""

    def __init__(self, ctx0, counts, group=None):
        self.ctx0, self.counts, self.group = ctx0, counts, group

    def gather(self, t, per_frame):
        cnt = [c * per_frame for c in self.counts]
        src = t.transpose(0, 1).contiguous()
        out = torch.empty((sum(cnt),) + tuple(src.shape[1:]), dtype=t.dtype)
        D.allgather_rows(out, cnt, src, self.group)
        return out.transpose(0, 1)


def _sp_problem(T=5, B=2, h=16, w=16):
    from oracle import seva_oracle as O
    from oracle import synth
    from seva_b200.model import Seva, SevaParams

    P = SevaParams(model_channels=64, num_frames=T)
    with torch.device("meta"):
        m = Seva(P)
    sd = synth.fill_state_dict({k: tuple(v.shape) for k, v in m.state_dict().items()}, seed=0)
    g = torch.Generator().manual_seed(5)
    n = B * T
    x = torch.randn(n, P.in_channels, h, w, generator=g)
    t = torch.full((n,), 499, dtype=torch.int64)
    y = torch.randn(n, 1, P.context_dim, generator=g)  # a different token per frame: context[::T] must be frame 0's
    dense = torch.randn(n, 6, h, w, generator=g)
    return O, P, sd, x, t, y, dense


def _worker(rank, world, port, ret):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        out = {}
        # ragged all-gather (frame partition 3 + 2)
        counts = [b - a for a, b in D.partition_frames(5, world)]
        mine = torch.arange(counts[rank] * 3, dtype=torch.float32).view(counts[rank], 3) + 100 * rank
        full = torch.empty(sum(counts), 3)
        D.allgather_rows(full, counts, mine)
        out["ragged"] = full
        # equal all-gather
        eq = torch.empty(4, 2)
        D.allgather_rows(eq, [2, 2], torch.full((2, 2), float(rank)))
        out["equal"] = eq
        # chunk data-parallel gather: 5 chunks over 2 ranks
        local = {k: torch.full((2, 3), float(k)) for k in D.schedule_chunks(5, world)[rank]}
        out["chunks"] = torch.stack(D.gather_chunks(local, 5))
        # CFG split through the sampler
        topo = D.Topology(world, rank, cfg=2, sp=1)
        groups = D.build_groups(topo)
        x0, c, uc, c2w, K, mask = _sampling_inputs()
        guider = D.CFGSplit(S.MultiviewCFG(1.2), groups["cfg"], half=topo.half)
        out["cfg_split"] = _run_sampler(guider, x0, c, uc, c2w, K, mask)
        # frame sequence-parallel restatement: each rank owns 3 / 2 of the 5 frames of both batch elements
        O, P, sd, x, t, y, dense = _sp_problem()
        T, B = 5, 2
        parts = D.partition_frames(T, world)
        t0, t1 = parts[rank]
        loc = lambda a: a.view(B, T, *a.shape[1:])[:, t0:t1].reshape(B * (t1 - t0), *a.shape[1:])
        sp = _OracleFrameGather(y[::T], [b - a for a, b in parts])
        with torch.inference_mode():
            o_loc = O.seva_forward(sd, P, loc(x), loc(t), loc(y), loc(dense), num_frames=t1 - t0, sp=sp)
        out["sp_local"] = o_loc.clone()
        ret[rank] = out
    finally:
        dist.destroy_process_group()


@pytest.fixture(scope="module")
def two_ranks():
    ctx = mp.get_context("spawn")
    with ctx.Manager() as mgr:
        ret = mgr.dict()
        port = _free_port()
        procs = [ctx.Process(target=_worker, args=(r, 2, port, ret)) for r in range(2)]
        for p in procs:
            p.start()
        for p in procs:
            p.join(600)   # a cold `import torch` in a spawned worker alone can take a minute on a busy box
            assert p.exitcode == 0, f"worker exit code {p.exitcode}"
        yield {r: dict(ret[r]) for r in range(2)}


def test_allgather_rows(two_ranks):
    exp = torch.cat([torch.arange(9, dtype=torch.float32).view(3, 3), torch.arange(6, dtype=torch.float32).view(2, 3) + 100])
    for r in range(2):
        assert torch.equal(two_ranks[r]["ragged"], exp)
        assert torch.equal(two_ranks[r]["equal"], torch.tensor([[0., 0.], [0., 0.], [1., 1.], [1., 1.]]))


def test_gather_chunks(two_ranks):
    exp = torch.stack([torch.full((2, 3), float(k)) for k in range(5)])
    for r in range(2):
        assert torch.equal(two_ranks[r]["chunks"], exp)


def test_cfg_split_matches_single_process(two_ranks):
    """uncond on rank 0, cond on rank 1: bit-identical to the one-process [uncond; cond] batch, on both ranks."""
    x0, c, uc, c2w, K, mask = _sampling_inputs()
    ref = _run_sampler(S.MultiviewCFG(1.2), x0, c, uc, c2w, K, mask)
    for r in range(2):
        assert torch.equal(two_ranks[r]["cfg_split"], ref)


def test_frame_sequence_parallel_restatement(two_ranks):
    """Frames 0-2 on rank 0, 3-4 on rank 1, K/V exchanged only at the 3D and time-mix self-attention: equals the
    one-process forward (fp32, reduction order inside SDPA differs -> 1e-4)."""
    O, P, sd, x, t, y, dense = _sp_problem()
    with torch.inference_mode():
        ref = O.seva_forward(sd, P, x, t, y, dense, num_frames=5)
    ref = ref.view(2, 5, *ref.shape[1:])
    parts = D.partition_frames(5, 2)
    for r, (t0, t1) in enumerate(parts):
        got = two_ranks[r]["sp_local"].view(2, t1 - t0, *ref.shape[2:])
        err = ((got - ref[:, t0:t1]).norm() / ref[:, t0:t1].norm()).item()
        assert err < 1e-4, (r, err)
