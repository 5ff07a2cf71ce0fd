"""Multi-GPU partitioning of the sampling path: one process per GPU over ``torch.distributed``.

Three axes, all optional and composable (world = chunks x cfg x frames):

* **chunk data-parallel** -- the second-pass chunks of the two-pass procedure (and every one-pass ``nearest*`` /
  ``interp*`` chunk) read only ``traj_prior_*`` / ``test_*`` tensors (seva/eval.py:1890-1906), never each other's
  samples: chunk k runs on rank k mod P with replicated weights, no data-path collective; the finished latents are
  gathered once (``gather_chunks``) and re-ordered like ``np.argsort(all_test_inds)`` (eval.py:1957-1959).
* **CFG split** -- ``VanillaCFG.prepare_inputs`` stacks [uncond; cond] into one batch (seva/sampling.py:231-242) and
  the halves never interact before the guidance combine (:213).  ``CFGSplit`` wraps any guider: rank ``half`` 0
  denoises the uncond half, rank 1 the cond half, one all-gather of the [T,4,h,w] fp32 result per step, after which
  every rank applies the identical combine + Euler update, so ``x`` stays replicated bit for bit.
* **frame sequence-parallel** -- every per-frame op (convs, GroupNorm, 2D attention, Linear/FF/LayerNorm) is local to
  the rank that owns the frame; the cross-frame ops (3D attention, time-mix attention, ``context[::T]``) exchange K/V
  between the ranks of the frame group (engine.py, ``FrameShard``).  ``partition_frames`` is the split both sides use.

Everything here is host logic over ``torch.distributed`` (NCCL on GPUs; the gloo tests run the same code on CPU
tensors).  Ranks must draw identical noise: seed every rank's generator identically before sampling.
"""
from __future__ import annotations

from dataclasses import dataclass

import torch
import torch.distributed as dist

_CAT_KEYS = ("vector", "crossattn", "concat", "replace", "dense_vector")  # seva/sampling.py:234


# ----------------------------------------------------------------------------------------------------
# integer bookkeeping (bit-exact, unit-tested on CPU)
# ----------------------------------------------------------------------------------------------------
def partition_frames(T: int, P: int) -> list[tuple[int, int]]:
    """Contiguous frame ranges [t0, t1) for P ranks; the first T % P ranks own one extra frame (T=21, P=4 ->
    6,5,5,5).  Ranks beyond T own an empty range."""
    if T <= 0 or P <= 0:
        raise ValueError(f"partition_frames needs T > 0 and P > 0 (got T={T}, P={P})")
    base, extra = divmod(T, P)
    out, t = [], 0
    for r in range(P):
        n = base + (1 if r < extra else 0)
        out.append((t, t + n))
        t += n
    return out


def schedule_chunks(n_chunks: int, world: int) -> list[list[int]]:
    """Round-robin chunk -> rank map (chunk k on rank k mod world): per-rank chunk id lists."""
    if n_chunks < 0 or world <= 0:
        raise ValueError(f"schedule_chunks needs n_chunks >= 0 and world > 0 (got {n_chunks}, {world})")
    return [list(range(r, n_chunks, world)) for r in range(world)]


@dataclass(frozen=True)
class Topology:
    """Rank layout world = chunks x cfg x frames, frames fastest: rank = (chunk_rank * cfg + half) * sp + frame_rank."""
    world: int
    rank: int
    cfg: int = 1     # 1 or 2
    sp: int = 1      # ranks sharing one chunk's frames

    def __post_init__(self):
        if self.cfg not in (1, 2):
            raise ValueError(f"cfg must be 1 or 2 (got {self.cfg})")
        if self.sp < 1 or self.world % (self.cfg * self.sp):
            raise ValueError(f"world {self.world} is not a multiple of cfg*sp = {self.cfg * self.sp}")
        if not 0 <= self.rank < self.world:
            raise ValueError(f"rank {self.rank} outside world {self.world}")

    @property
    def frame_rank(self) -> int:
        return self.rank % self.sp

    @property
    def half(self) -> int:
        return (self.rank // self.sp) % self.cfg

    @property
    def chunk_rank(self) -> int:
        return self.rank // (self.sp * self.cfg)

    @property
    def chunk_world(self) -> int:
        return self.world // (self.sp * self.cfg)

    def frame_group_ranks(self) -> list[int]:
        base = self.rank - self.frame_rank
        return list(range(base, base + self.sp))

    def cfg_group_ranks(self) -> list[int]:
        base = self.chunk_rank * self.cfg * self.sp + self.frame_rank
        return [base + h * self.sp for h in range(self.cfg)]

    def model_group_ranks(self) -> list[int]:
        """All ranks that cooperate on one chunk (cfg x frames)."""
        base = self.chunk_rank * self.cfg * self.sp
        return list(range(base, base + self.cfg * self.sp))


def build_groups(topo: Topology) -> dict:
    """Create every sub-group on every rank (``new_group`` is collective) and return this rank's
    {"frames": group, "cfg": group, "model": group}; a size-1 axis maps to None."""
    mine: dict = {"frames": None, "cfg": None, "model": None}
    for r in range(topo.world):
        t = Topology(topo.world, r, topo.cfg, topo.sp)
        for name, ranks in (("frames", t.frame_group_ranks()), ("cfg", t.cfg_group_ranks()),
                            ("model", t.model_group_ranks())):
            if len(ranks) > 1 and ranks[0] == r:  # the lowest rank of each group names it once
                g = dist.new_group(ranks)
                if topo.rank in ranks:
                    mine[name] = g
    return mine


# ----------------------------------------------------------------------------------------------------
# collectives
# ----------------------------------------------------------------------------------------------------
def allgather_rows(out: torch.Tensor, counts: list[int], inp: torch.Tensor, group=None) -> torch.Tensor:
    """Variable-size all-gather along dim 0: rank r contributes ``inp`` (counts[r] rows), every rank ends with all
    rows in rank order in ``out`` (sum(counts) rows).  Equal counts use one all-gather; ragged counts (T=21 frames
    over 2/4/8 ranks) use one broadcast per contributing rank inside the group -- the same wire traffic."""
    P = dist.get_world_size(group)
    if len(counts) != P:
        raise ValueError(f"allgather_rows: {len(counts)} counts for a group of {P}")
    if out.shape[0] != sum(counts) or inp.shape[0] != counts[dist.get_rank(group)]:
        raise ValueError("allgather_rows: buffer sizes do not match counts")
    if len(set(counts)) == 1:
        dist.all_gather_into_tensor(out, inp.contiguous(), group=group)
        return out
    r0 = 0
    me = dist.get_rank(group)
    for r, n in enumerate(counts):
        if n:
            piece = out[r0:r0 + n]
            if r == me:
                piece.copy_(inp)
            dist.broadcast(piece, src=dist.get_global_rank(group, r) if group is not None else r, group=group)
        r0 += n
    return out


def gather_chunks(local: dict, n_chunks: int, group=None) -> list:
    """Collect the per-chunk results (chunk id -> tensor, all of one shape/dtype) of a ``schedule_chunks`` run.
    Every rank returns the full list in chunk order (ranks that own fewer chunks contribute a zero pad)."""
    P = dist.get_world_size(group)
    me = dist.get_rank(group)
    mine = schedule_chunks(n_chunks, P)[me]
    if sorted(local) != mine:
        raise ValueError(f"rank {me} holds chunks {sorted(local)}, expected {mine}")
    rounds = (n_chunks + P - 1) // P
    proto = next(iter(local.values())) if local else None
    # shapes travel as objects (host metadata only); ranks with no chunk learn the shape from rank 0
    meta = [None] * P
    dist.all_gather_object(meta, None if proto is None else (tuple(proto.shape), str(proto.dtype), str(proto.device)),
                           group=group)
    ref = next((m for m in meta if m is not None), None)
    if ref is None:
        return []
    dtype = getattr(torch, ref[1].split(".")[1])
    device = proto.device if proto is not None else torch.device(ref[2] if not ref[2].startswith("cuda")
                                                                 else f"cuda:{torch.cuda.current_device()}")
    out: list = [None] * n_chunks
    for i in range(rounds):
        k = i * P + me
        send = local[k] if k in local else torch.zeros(ref[0], dtype=dtype, device=device)
        flat = torch.empty((P * send.numel(),), dtype=dtype, device=device)  # dim-0 concatenation (gloo and NCCL)
        dist.all_gather_into_tensor(flat, send.contiguous().view(-1), group=group)
        buf = flat.view((P,) + tuple(ref[0]))
        for r in range(P):
            if i * P + r < n_chunks:
                out[i * P + r] = buf[r]
    return out


# ----------------------------------------------------------------------------------------------------
# CFG split
# ----------------------------------------------------------------------------------------------------
class CFGSplit(object):
    """Guider wrapper that places the uncond and cond halves of classifier-free guidance on two ranks.

    ``prepare_inputs`` (seva/sampling.py:231-242) returns only this rank's half -- uc[k] on ``half`` 0, c[k] on
    ``half`` 1 -- so the denoiser runs a batch of T instead of 2T; ``gather`` restores the reference's
    [uncond; cond] stacking with one all-gather, and ``__call__`` then applies the wrapped guider unchanged.
    Drop-in: ``EulerEDMSampler(discretization, CFGSplit(MultiviewCFG(cfg_min), group), ...)``.
    """

    def __init__(self, guider, group=None, half: int | None = None):
        self.guider = guider
        self.group = group
        size = dist.get_world_size(group)
        if size != 2:
            raise ValueError(f"CFGSplit needs a group of exactly 2 ranks (got {size})")
        self.half = dist.get_rank(group) if half is None else half

    def __getattr__(self, name):  # scale_rule / scale_schedule / guidance / scale_min of the wrapped guider
        return getattr(self.__dict__["guider"], name)

    def prepare_inputs(self, x: torch.Tensor, s: torch.Tensor, c: dict, uc: dict):
        from .sampling import COND_EPOCH, _EPOCHS, _same_cond, _version

        pick = uc if self.half == 0 else c
        ident = [(k, pick[k], pick[k], _version(pick[k]), 0) for k in c if k in _CAT_KEYS]
        cache = self.__dict__.get("_cond_cache")
        if cache is None or not _same_cond(cache[0], ident):  # one conditioning epoch per (half, loop), as VanillaCFG
            cache = self.__dict__["_cond_cache"] = (ident, next(_EPOCHS))
        c_out = {COND_EPOCH: cache[1]}
        for k in c:
            if k in _CAT_KEYS:
                c_out[k] = pick[k]
            else:
                assert c[k] == uc[k]
                c_out[k] = c[k]
        return x, s, c_out

    def reset_cache(self) -> None:
        self.__dict__.pop("_cond_cache", None)
        reset = getattr(self.guider, "reset_cache", None)
        if reset is not None:
            reset()

    def gather(self, denoised: torch.Tensor) -> torch.Tensor:
        out = torch.empty((2 * denoised.shape[0],) + tuple(denoised.shape[1:]), dtype=denoised.dtype,
                          device=denoised.device)
        dist.all_gather_into_tensor(out, denoised.contiguous(), group=self.group)
        return out

    def __call__(self, x: torch.Tensor, sigma, scale, **guider_kwargs) -> torch.Tensor:
        """``x`` is this rank's half; the halves are gathered, then combined by the wrapped guider."""
        return self.guider(self.gather(x), sigma, scale, **guider_kwargs)

    def frame_scale(self, x, sigma, scale, **guider_kwargs):
        fs = getattr(self.guider, "frame_scale", None)
        return None if fs is None else fs(x, sigma, scale, **guider_kwargs)
