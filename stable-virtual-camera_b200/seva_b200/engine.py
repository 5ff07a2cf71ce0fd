"""Execution plan for one ``Seva.forward`` shape: packed weights, static workspaces, and a flat list of C-ABI
kernel launches (see DESIGN.md "Data layout in HBM").

Layouts
  * "padded": bf16 [N*(H+2)*(W+2), C] channels-last flat image with a zero border (ResBlock / conv land); a 3x3
    conv is nine row-shifted GEMM taps over it.
  * "tokens": bf16 [N*H*W, C] (transformer land); "(b t)(h w) c -> b (t h w) c" (transformer.py:237) is a view.
  * skip concats ``torch.cat([h, hs.pop()], 1)`` (model.py:207) are never materialised: both producers write
    column slices of one buffer (row stride = C_h + C_skip).
All buffers are allocated once per (N, T, h, w) and all launch descriptors are built once, so ``run()`` is a
loop of foreign calls with constant arguments (and is CUDA-graph capturable).
"""
from __future__ import annotations

import ctypes as C
import math
import os

import torch

from . import _lib, ops, pack
from ._lib import SevaB200Error
from .model import (Downsample, MultiviewTransformer, ResBlock, Seva, Upsample)

BF16, F32 = torch.bfloat16, torch.float32


# ----------------------------------------------------------------------------------------------------
# weight packing
# ----------------------------------------------------------------------------------------------------
def _f32(p):
    return p.detach().float().contiguous()


def _pack_ff(ff, dev):
    w, b = ff.net[0].proj.weight, ff.net[0].proj.bias
    bn = ops.pick_bn(w.shape[0], True)
    wp, bp = pack.pack_geglu(w, b, bn)
    return dict(w1=wp, b1=bp, bn1=bn, w2=pack.pack_linear(ff.net[2].weight), b2=_f32(ff.net[2].bias))


def _no_launch(*_args) -> int:
    """placeholder plan entry (the timestep-table gather before its table exists)"""
    return 0


def _pad_head_rows(w, heads, dh):
    """[heads * dh, K] -> [heads * 64, K]: every head's rows followed by 64 - dh zero rows (dh <= 64)."""
    if dh == 64:
        return w
    out = w.new_zeros(heads, 64, w.shape[1])
    out[:, :dh] = w.reshape(heads, dh, w.shape[1])
    return out.reshape(heads * 64, w.shape[1])


def _pad_head_cols(w, heads, dh):
    """[N, heads * dh] -> [N, heads * 64] with zero columns at the padded head dimensions."""
    return w if dh == 64 else _pad_head_rows(w.t(), heads, dh).t()


def _pack_attn_self(at, heads=None, dh=64):
    hp = lambda w: _pad_head_rows(w.detach(), heads, dh)
    return dict(wqkv=pack.pack_linear(torch.cat([hp(at.to_q.weight), hp(at.to_k.weight), hp(at.to_v.weight)], 0)),
                wo=pack.pack_linear(_pad_head_cols(at.to_out[0].weight.detach(), heads, dh)), bo=_f32(at.to_out[0].bias))


def _pack_attn_cross(at, heads=None, dh=64):
    hp = lambda w: pack.pack_linear(_pad_head_rows(w.detach(), heads, dh))
    # wv1 / wo1: the unpadded pair of the single-context-token shortcut (to_out(to_v(ctx)), no head structure)
    return dict(wq=hp(at.to_q.weight), wk=hp(at.to_k.weight), wv=hp(at.to_v.weight),
                wo=pack.pack_linear(_pad_head_cols(at.to_out[0].weight.detach(), heads, dh)),
                wv1=pack.pack_linear(at.to_v.weight), wo1=pack.pack_linear(at.to_out[0].weight),
                bo=_f32(at.to_out[0].bias))


def _pack_ln(ln):
    return dict(g=_f32(ln.weight), b=_f32(ln.bias), eps=float(ln.eps))


# LayerNorm folded into the consuming GEMM's epilogue.  Opt-in: measured SLOWER in the step (r2b/r2c: the K = 320 GEGLU goes
# 468 -> 683 us, qkv 185 -> 283 us -- these GEMMs are epilogue-bound and the fold adds 2 FFMA + 1 LDS.128 per 4 values
# there, which costs more than the 24 us LayerNorm launch it removes).
LN_FOLD = os.environ.get("SEVA_LN_FOLD", "0") != "0"
FUSED_GN_STATS = os.environ.get("SEVA_GN_FUSED_STATS", "1") != "0"  # GroupNorm statistics from the producing GEMM


def _can_fold(C_: int) -> bool:
    return LN_FOLD and C_ % 32 == 0 and 0 < ops.ln_slots(C_) <= 10


def _fold_ln_linear(ln, w, b):
    """Linear(LayerNorm(x)) = rstd * (x . (W gamma)^T - mean * colsum) + (W beta + b):
    returns (W * gamma in bf16, colsum of the ROUNDED folded weight (what the tensor cores multiply), W beta + b)."""
    wf = w.detach().float()
    wg = (wf * ln.weight.detach().float()[None]).to(torch.bfloat16)
    bias2 = wf @ ln.bias.detach().float()
    if b is not None:
        bias2 = bias2 + b.detach().float()
    return wg.contiguous(), wg.float().sum(1).contiguous(), bias2.contiguous()


def _pack_ff_ln(ff, ln):
    """FeedForward whose input LayerNorm is folded into the GEGLU projection."""
    w, b = ff.net[0].proj.weight, ff.net[0].proj.bias
    bn = ops.pick_bn(w.shape[0], True)
    wg, colsum, bias2 = _fold_ln_linear(ln, w, b)
    wp, bp = pack.pack_geglu(wg, bias2, bn)
    _, cp = pack.pack_geglu(wg, colsum, bn)
    return dict(w1=wp, b1=bp, cs1=cp, bn1=bn, eps=float(ln.eps), w2=pack.pack_linear(ff.net[2].weight),
                b2=_f32(ff.net[2].bias))


def _pack_attn_self_ln(at, ln):
    wg, colsum, bias2 = _fold_ln_linear(ln, torch.cat([at.to_q.weight, at.to_k.weight, at.to_v.weight], 0), None)
    return dict(wqkv=wg, csqkv=colsum, bqkv=bias2, eps=float(ln.eps), wo=pack.pack_linear(at.to_out[0].weight),
                bo=_f32(at.to_out[0].bias))


def pack_weights(model: Seva) -> dict:
    """id(module) -> dict of kernel-layout tensors (bf16 K-major matrices, fp32 vectors)."""
    dev = next(model.parameters()).device
    P: dict = {}
    with torch.no_grad():
        te = model.time_embed
        P["time_embed"] = dict(w0=pack.pack_linear(te[0].weight), b0=_f32(te[0].bias),
                               w2=pack.pack_linear(te[2].weight), b2=_f32(te[2].bias))
        conv_in = model.input_blocks[0][0]
        P["conv_in"] = dict(w=pack.pack_conv3x3(conv_in.weight, cin_pad=64), b=_f32(conv_in.bias))
        P["out"] = dict(g=_f32(model.out[0].weight), b=_f32(model.out[0].bias), eps=float(model.out[0].eps),
                        w=pack.pack_conv3x3(model.out[2].weight), cb=_f32(model.out[2].bias))
        for m in model.modules():
            if isinstance(m, ResBlock):
                d = dict(
                    g1=_f32(m.in_layers[0].weight), b1=_f32(m.in_layers[0].bias), eps1=float(m.in_layers[0].eps),
                    w1=pack.pack_conv3x3(m.in_layers[2].weight), cb1=_f32(m.in_layers[2].bias),
                    we=pack.pack_linear(m.emb_layers[1].weight), be=_f32(m.emb_layers[1].bias),
                    wd=_f32(m.dense_emb_layers[0].weight.reshape(2 * m.channels, -1)),
                    bd=_f32(m.dense_emb_layers[0].bias),
                    g2=_f32(m.out_layers[0].weight), b2=_f32(m.out_layers[0].bias), eps2=float(m.out_layers[0].eps),
                    w2=pack.pack_conv3x3(m.out_layers[3].weight), cb2=_f32(m.out_layers[3].bias))
                if not isinstance(m.skip_connection, torch.nn.Identity):
                    d["ws"] = pack.pack_conv1x1(m.skip_connection.weight)
                    d["bs"] = _f32(m.skip_connection.bias)
                P[id(m)] = d
            elif isinstance(m, MultiviewTransformer):
                fold = _can_fold(m.in_channels) and m.d_head == 64
                hd = dict(heads=m.n_heads, dh=m.d_head)
                pairs = []
                for tb, tm in zip(m.transformer_blocks, m.time_mix_blocks):   # transformer_depth pairs (transformer.py:234)
                    pairs.append(dict(
                        s_attn1=_pack_attn_self_ln(tb.attn1, tb.norm1) if fold else _pack_attn_self(tb.attn1, **hd),
                        s_attn2=_pack_attn_cross(tb.attn2, **hd),
                        s_ff=_pack_ff_ln(tb.ff, tb.norm3) if fold else _pack_ff(tb.ff, dev),
                        s_n1=_pack_ln(tb.norm1), s_n2=_pack_ln(tb.norm2), s_n3=_pack_ln(tb.norm3),
                        t_nin=_pack_ln(tm.norm_in),
                        t_ffin=_pack_ff_ln(tm.ff_in, tm.norm_in) if fold else _pack_ff(tm.ff_in, dev),
                        t_attn1=_pack_attn_self_ln(tm.attn1, tm.norm1) if fold else _pack_attn_self(tm.attn1, **hd),
                        t_attn2=_pack_attn_cross(tm.attn2, **hd),
                        t_ff=_pack_ff_ln(tm.ff, tm.norm3) if fold else _pack_ff(tm.ff, dev),
                        t_n1=_pack_ln(tm.norm1), t_n2=_pack_ln(tm.norm2), t_n3=_pack_ln(tm.norm3)))
                P[id(m)] = dict(
                    gn=dict(g=_f32(m.norm.weight), b=_f32(m.norm.bias), eps=float(m.norm.eps)),
                    win=pack.pack_linear(m.proj_in.weight), bin=_f32(m.proj_in.bias),
                    wout=pack.pack_linear(m.proj_out.weight), bout=_f32(m.proj_out.bias), pairs=pairs, ln_fold=fold)
            elif isinstance(m, Downsample):
                P[id(m)] = dict(w=pack.pack_conv3x3(m.op.weight), b=_f32(m.op.bias))
            elif isinstance(m, Upsample):
                P[id(m)] = dict(wph=pack.pack_upconv_phases(m.conv.weight), b=_f32(m.conv.bias))
    return P


# ----------------------------------------------------------------------------------------------------
# frame sequence-parallel shard: this rank owns frames [t0, t1) of every batch element
# ----------------------------------------------------------------------------------------------------
class FrameShard:
    def __init__(self, group, T_total: int, size: int | None = None, rank: int | None = None):
        """``size`` / ``rank`` default to the process group's; passing them builds the shard of another rank for the
        one-process emulation of the frame group (tests/test_model_gpu.py), whose ``gather`` must not be called."""
        from .dist import partition_frames

        self.group = group
        if size is None:
            import torch.distributed as dist

            size, rank = dist.get_world_size(group), dist.get_rank(group)
        self.size, self.rank = size, rank
        self.T_total = T_total
        self.parts = partition_frames(T_total, self.size)
        self.t0, self.t1 = self.parts[self.rank]
        self.counts = [b - a for a, b in self.parts]
        if min(self.counts) < 1:
            raise SevaB200Error(f"{T_total} frames cannot be split over {self.size} ranks (a rank would own none)")
        self._stage: dict = {}   # padded all-gather staging buffers per (B, rows, C)

    @property
    def T_local(self) -> int:
        return self.t1 - self.t0

    def space_parts(self, S: int) -> list:
        """Contiguous ranges of the S spatial positions, one per rank (time-mix all-to-all)."""
        from .dist import partition_frames

        return partition_frames(S, self.size)

    def stage_flat(self, n_send: int, n_recv: int, dtype, device):
        key = ("flat", n_send, n_recv, dtype, device)
        st = self._stage.get(key)
        if st is None:
            st = (torch.empty(n_send, dtype=dtype, device=device), torch.empty(n_recv, dtype=dtype, device=device))
            self._stage[key] = st
        return st

    def local(self, x: torch.Tensor, B: int) -> torch.Tensor:
        """[B*T_total, ...] -> this rank's [B*T_local, ...] (contiguous copy)."""
        return x.reshape(B, self.T_total, *x.shape[1:])[:, self.t0:self.t1].reshape(B * self.T_local, *x.shape[1:]) \
            .contiguous()

    def gather(self, out: torch.Tensor, loc: torch.Tensor, B: int, rows_per_frame: int) -> None:
        """out [B*T_total*rpf, C] <- every rank's loc [B*T_r*rpf, C], frames in global order per batch element."""
        import torch.distributed as dist

        nl, nf = self.T_local * rows_per_frame, self.T_total * rows_per_frame
        if B == 1 and len(set(self.counts)) == 1:
            dist.all_gather_into_tensor(out, loc, group=self.group)
            return
        # ONE collective per call whatever the split: every rank contributes a block padded to the largest share
        # ([B, cmax * rpf, C]); the P blocks are then scattered to their frame ranges with one strided copy each.
        # (A broadcast per rank and batch element -- 2 P B launches per site, 23 sites per forward -- cost more
        # than the attention it feeds at 8 GPUs.)
        cmax = max(self.counts) * rows_per_frame
        C_ = loc.shape[1]
        key = (B, cmax, C_, loc.dtype, loc.device)
        st = self._stage.get(key)
        if st is None:
            st = (torch.zeros(B, cmax, C_, dtype=loc.dtype, device=loc.device),
                  torch.empty(self.size, B, cmax, C_, dtype=loc.dtype, device=loc.device))
            self._stage[key] = st
        send, recv = st
        send[:, :nl].copy_(loc.view(B, nl, C_))
        dist.all_gather_into_tensor(recv, send, group=self.group)
        o = out.view(B, nf, C_)
        for r, (a, b) in enumerate(self.parts):
            o[:, a * rows_per_frame:b * rows_per_frame].copy_(recv[r, :, :(b - a) * rows_per_frame])


class _FrameGather:
    """Plan entry (called like a C-ABI launch): K/V of all frames <- the frame group."""
    __name__ = "frame_gather_kv"

    def __init__(self, shard: FrameShard, out, loc, B, rows_per_frame):
        self.a = (shard, out, loc, B, rows_per_frame)

    def __call__(self, stream) -> int:
        shard, out, loc, B, rpf = self.a
        shard.gather(out, loc, B, rpf)
        return 0


class _RingAttention:
    """Plan entry (called like a C-ABI launch): frame sequence-parallel 3D self-attention with the K/V exchange
    OVERLAPPED with the attention itself (SURVEY 8e "ring variant").

    Every rank holds q and [k | v] of its own frames.  The peers' [k | v] blocks travel on a second stream -- step s of
    P - 1: send mine to rank + s, receive the block of rank - s (NCCL point-to-point over NVSwitch, straight into a
    per-peer buffer of exactly that peer's size: no padded staging, no scatter copies) -- while the compute stream
    attends to the blocks already here, local block first: one ``seva_attn_partial`` launch per block (fp32 result +
    log-sum-exp of its scores), then ``seva_attn_merge`` (online-softmax merge of the P partial results).  Same
    mathematics as one attention over the gathered keys; the exchange is hidden behind P - 1 of the P block attentions.
    ``mode`` = "allgather": one all-gather on the second stream instead of the P - 1 steps (overlaps the local block only).
    """
    __name__ = "ring_attn3d"

    def __init__(self, lib, shard: FrameShard, q, kvl, o, B, heads, S, C_, mode="ring", scale=0.125):
        self.lib, self.shard, self.q, self.kvl, self.o = lib, shard, q, kvl, o
        self.B, self.heads, self.S, self.C = B, heads, S, C_
        self.mode, self.scale = mode, float(scale)   # scale: 1 / sqrt(true head width), narrow heads run zero-padded
        dev, P = q.device, shard.size
        self.Lq = shard.T_local * S
        self.parts = torch.empty(P, B, self.Lq, C_, device=dev, dtype=F32)
        self.lse = torch.empty(P, B, heads, self.Lq, device=dev, dtype=F32)
        self.peer = [None if r == shard.rank else torch.empty(B, shard.counts[r] * S, 2 * C_, device=dev, dtype=BF16)
                     for r in range(P)]
        self.n_kernels = P + 1
        self._events = None
        st = shard._stage.get("ring_stream")
        if st is None:
            st = shard._stage["ring_stream"] = torch.cuda.Stream(device=dev)
        self.comm = st

    def start(self, stream=None) -> int:
        """Enqueue the exchange on the second stream (plan entry placed right after the K/V projection, so that the
        blocks travel while q is still being projected)."""
        import torch.distributed as dist

        sh, P, me = self.shard, self.shard.size, self.shard.rank
        cur = torch.cuda.current_stream()
        ready = torch.cuda.Event()
        ready.record(cur)                      # kvl was produced on the compute stream
        events = {}
        with torch.cuda.stream(self.comm):
            self.comm.wait_event(ready)
            if self.mode == "ring":
                ranks = dist.get_process_group_ranks(sh.group)
                for step in range(1, P):
                    dst, src = (me + step) % P, (me - step) % P
                    ops_ = [dist.P2POp(dist.isend, self.kvl, ranks[dst], group=sh.group),
                            dist.P2POp(dist.irecv, self.peer[src], ranks[src], group=sh.group)]
                    for req in dist.batch_isend_irecv(ops_):
                        req.wait()
                    events[src] = torch.cuda.Event()
                    events[src].record(self.comm)
            else:
                dist.all_gather_into_tensor(self._gathered(), self._send_padded(), group=sh.group)
                ev = torch.cuda.Event()
                ev.record(self.comm)
                events = {r: ev for r in range(P) if r != me}
        self._events = events
        return 0

    def block(self, r):
        """(tensor, keys, batch stride) of rank r's [k | v] block as this rank holds it."""
        sh = self.shard
        if r == sh.rank:
            return self.kvl, self.Lq, self.Lq * 2 * self.C
        if self.mode == "ring":
            return self.peer[r], sh.counts[r] * self.S, sh.counts[r] * self.S * 2 * self.C
        kv = self._gathered()[r]
        return kv, sh.counts[r] * self.S, kv.stride(0)

    def compute(self, stream, events=None) -> int:
        """Local block first, then the peers' blocks in arrival order, then the merge."""
        P, me = self.shard.size, self.shard.rank
        cur = torch.cuda.current_stream()
        for r in [me] + [(me - step) % P for step in range(1, P)]:
            if events is not None and r != me:
                cur.wait_event(events[r])
            kv, Lk, bs = self.block(r)
            rc = self.lib.seva_attn_partial(self.q.data_ptr(), kv.data_ptr(), kv.data_ptr() + 2 * self.C,
                                            self.parts[r].data_ptr(), self.lse[r].data_ptr(), self.B, self.heads, self.Lq,
                                            Lk, self.C, 2 * self.C, 2 * self.C, self.Lq * self.C, bs, bs, self.scale, stream)
            if rc:
                return rc
        return self.lib.seva_attn_merge(self.parts.data_ptr(), self.lse.data_ptr(), self.o.data_ptr(), P, self.B, self.heads,
                                        self.Lq, self.o.stride(0), self.Lq * self.o.stride(0), stream)

    def __call__(self, stream) -> int:
        if self._events is None:
            self.start(stream)
        events, self._events = self._events, None
        return self.compute(stream, events)

    # ---- all-gather variant: padded blocks (the frame split is uneven: T = 21 never divides) -----------------------
    def _gathered(self):
        sh = self.shard
        key = ("ring_ag", self.B, max(sh.counts) * self.S, self.C)
        st = sh._stage.get(key)
        if st is None:
            cmax = max(sh.counts) * self.S
            st = sh._stage[key] = (torch.zeros(self.B, cmax, 2 * self.C, device=self.q.device, dtype=BF16),
                                   torch.empty(sh.size, self.B, cmax, 2 * self.C, device=self.q.device, dtype=BF16))
        return st[1]

    def _send_padded(self):
        self._gathered()
        send = self.shard._stage[("ring_ag", self.B, max(self.shard.counts) * self.S, self.C)][0]
        send[:, : self.Lq].copy_(self.kvl.view(self.B, self.Lq, 2 * self.C))
        return send


class _RingStart:
    """Plan entry: begin a ``_RingAttention``'s K/V exchange (no kernel of ours)."""
    __name__ = "ring_exchange_start"

    def __init__(self, ring: _RingAttention):
        self.ring = ring

    def __call__(self, stream) -> int:
        return self.ring.start(stream)


class _SpaceExchange:
    """Plan entry (called like a C-ABI launch): all-to-all between the frame-sharded token layout
    [B, T_local, S, C] and the space-sharded layout [B, T_total, S_local, C].

    Time-mix attention needs all FRAMES of one spatial position, not all positions: instead of gathering K/V of
    every frame on every rank (2C per token of the whole chunk, P ranks -> (P-1)/P * 2C * P * local tokens received),
    the packed q|k|v of the local frames is transposed to "all frames of my slice of positions" (3C per local token
    sent once) and the attention output (C per token) is transposed back: 4C against 2C*P per local token, i.e. 4x less
    NVLink traffic at 8 ranks, and the unsharded ``seva_timemix_attn`` kernel runs unchanged on the slice."""
    __name__ = "frame_space_a2a"

    def __init__(self, shard: FrameShard, to_space: bool, src, dst, B: int, S: int):
        self.shard, self.to_space, self.src, self.dst, self.B, self.S = shard, to_space, src, dst, B, S
        self.sparts = shard.space_parts(S)

    def _frame_sharded(self, t):   # per peer d: the rows of MY frames at peer d's positions
        sh = self.shard
        v = t.view(self.B, sh.T_local, self.S, t.shape[1])
        return [v[:, :, a:b] for a, b in self.sparts]

    def _space_sharded(self, t):   # per peer r: the rows of peer r's frames at MY positions
        sh = self.shard
        a, b = self.sparts[sh.rank]
        v = t.view(self.B, sh.T_total, b - a, t.shape[1])
        return [v[:, t0:t1] for t0, t1 in sh.parts]

    def views_out(self):
        return self._frame_sharded(self.src) if self.to_space else self._space_sharded(self.src)

    def views_in(self):
        return self._space_sharded(self.dst) if self.to_space else self._frame_sharded(self.dst)

    def __call__(self, stream) -> int:
        import torch.distributed as dist

        outs, ins = self.views_out(), self.views_in()
        send, recv = self.shard.stage_flat(self.src.numel(), self.dst.numel(), self.src.dtype, self.src.device)
        o = 0
        for v in outs:
            send[o:o + v.numel()].view(v.shape).copy_(v)
            o += v.numel()
        dist.all_to_all_single(recv, send, [v.numel() for v in ins], [v.numel() for v in outs], group=self.shard.group)
        o = 0
        for v in ins:
            v.copy_(recv[o:o + v.numel()].view(v.shape))
            o += v.numel()
        return 0


# ----------------------------------------------------------------------------------------------------
# plan
# ----------------------------------------------------------------------------------------------------
class Engine:
    def __init__(self, model: Seva, packed: dict, n: int, num_frames: int, h: int, w: int, ctx_len: int,
                 shard: FrameShard | None = None):
        p = model.params
        nlev = len(p.channel_mult)
        if h % (1 << (nlev - 1)) or w % (1 << (nlev - 1)):
            raise SevaB200Error(f"latent size {h}x{w} must be divisible by {1 << (nlev - 1)}")
        self.lib = _lib.load()
        self.dev = next(model.parameters()).device
        self.P = packed
        # with a frame shard, n / num_frames are LOCAL (this rank's frames); Tk = frames the cross-frame ops see
        self.N, self.T, self.B = n, num_frames, n // num_frames
        self.shard = shard
        self.Tk = num_frames if shard is None else shard.T_total
        if shard is not None and shard.T_local != num_frames:
            raise SevaB200Error("frame shard / engine frame count mismatch")
        self.h, self.w, self.L = h, w, ctx_len
        self.ctx_dim = p.context_dim
        self.unflatten = set(p.unflatten_names)
        self.HW = [(h >> l, w >> l) for l in range(nlev)]
        self.calls: list = []
        self.meta: list = []
        self.keep: list = []   # ctypes structs / tensors referenced by raw pointer
        self.bufs: dict = {}
        self.n_launch = 0
        self.gemv_groups: dict = {}
        self.gn_src: dict = {}   # (data_ptr, columns) of a GEMM output -> (partial-statistics buffer, rows per image)
        # time-mix attention under a frame shard: all-to-all to a space-sharded layout (default) or K/V all-gather
        # (measured at T=21: 8 ranks 26.4 vs 30.2 ms/step in favour of the all-to-all, 4 ranks a tie, 2 ranks 59.0 vs 56.7)
        env = os.environ.get("SEVA_TIMEMIX_A2A")
        self.tm_a2a = (shard is not None and shard.size > 4) if env is None else env != "0"
        self.walk: dict = {}    # storage -> direction (True = descending rows) of the last kernel that touched it
        self.zigzag = os.environ.get("SEVA_ZIGZAG", "0") == "1"  # measured: no gain at T=21
        self.timing_sink: list | None = None  # set by bench.py to collect per-launch CUDA events
        # Step-invariant hoisting: launches that read only the conditioning (Plucker resizes, Lk = 1
        # cross-attention vectors) are skipped while ``bind`` is handed the same conditioning epoch as last time.
        self._invariant: set = set()
        # launches that depend on the timestep only (t-emb, time_embed, emb_layers): with the per-index table (N2) a step
        # runs ONE gather instead; see _build_emb_table
        self._emb_calls: set = set()
        self._emb_outs: list = []
        self._emb_gather_idx = None
        self._emb_table = None
        self._emb_rows = 0
        self._use_emb_table = False
        self.emb_table_enabled = os.environ.get("SEVA_EMB_TABLE", "1") != "0"
        self._epoch = None
        self._skip_invariant = False
        # The plan is static (constant pointers and sizes), so after one eager run per variant (all launches / the
        # conditioning-only launches skipped) it is captured into a CUDA graph and replayed: one driver call per forward
        # instead of ~600 ctypes calls, each with its cuTensorMapEncodeTiled (the descriptors are baked into the
        # captured kernel parameters).  Not used while per-launch timing is on; SEVA_CUDA_GRAPH=0 turns it off.  With a
        # frame shard the plan holds torch.distributed collectives: capturing those is opt-in (SEVA_SHARD_GRAPH=1, every
        # rank alike).
        self.use_graph = os.environ.get("SEVA_CUDA_GRAPH", "1") != "0" and (
            shard is None or os.environ.get("SEVA_SHARD_GRAPH", "0") == "1")
        self._graphs: dict = {}
        self._eager_runs: dict = {}

        # static inputs
        self.net_in = self.buf("net_in", self.rows_p(0), 64)
        self.t_in = torch.zeros(n, device=self.dev, dtype=torch.int64)
        self.ctx_in = torch.zeros(n, ctx_len, p.context_dim, device=self.dev, dtype=F32)
        self.ctx0_in = torch.zeros(self.B, ctx_len, p.context_dim, device=self.dev, dtype=F32)
        self.dense_in = torch.zeros(n, p.dense_in_channels, h, w, device=self.dev, dtype=F32)
        self.eps = torch.zeros(self.rows_p(0), p.out_channels, device=self.dev, dtype=F32)
        self._build(model)

    # ---- helpers -------------------------------------------------------------------------------
    def rows_p(self, l):
        H, W = self.HW[l]
        return self.N * (H + 2) * (W + 2)

    def rows_c(self, l):
        H, W = self.HW[l]
        return self.N * H * W

    def buf(self, name, rows, cols, dtype=BF16):
        key = (name, rows, cols, dtype)
        t = self.bufs.get(key)
        if t is None:
            t = torch.zeros(rows, cols, device=self.dev, dtype=dtype)
            self.bufs[key] = t
        return t

    # L2-aware traversal (include/seva_b200.h, SevaGemm.reverse): a kernel walks its main input in the opposite
    # row order to the kernel that last touched that buffer, so the most recently written / read ~100 MB are found
    # in the 126 MB L2 instead of DRAM.  ``walk`` tracks the last direction per underlying storage.
    def _dir(self, src, *touched) -> bool:
        key = src.untyped_storage().data_ptr()
        rev = (not self.walk.get(key, False)) if self.zigzag else False
        for t in (src,) + touched:
            if t is not None:
                self.walk[t.untyped_storage().data_ptr()] = rev
        return rev

    def _touch_forward(self, *tensors) -> None:
        for t in tensors:
            if t is not None:
                self.walk[t.untyped_storage().data_ptr()] = False

    def add(self, fn, *args, flops: float = 0.0, nbytes: float = 0.0, tag: str = ""):
        """Append one kernel launch; flops / nbytes are the ALGORITHMIC work of the reference op it replaces
        (2*MAC for Linear/conv on unpadded pixels, 4*B*H*Lq*Lk*D for attention, activation bytes for norms)."""
        self.calls.append((fn, args))
        self.meta.append((fn.__name__, tag, float(flops), float(nbytes)))
        if isinstance(fn, _RingAttention):
            self.n_launch += fn.n_kernels
        elif not isinstance(fn, (_FrameGather, _SpaceExchange, _RingStart)):  # collectives are NCCL's kernels, not ours
            self.n_launch += 1

    # ---- op emitters (each appends exactly one kernel launch unless noted) ---------------------
    def gemm(self, A, W, out, *, bias=None, M=None, ntaps=1, tap_wp=0, rowbias=None, rowbias_rpi=0, residual=None,
             out_fp32=False, geglu=False, out_map=0, H=0, Wd=0, mask_border=False, Hp=0, Wp=0, bn=0, Kc=None,
             alg_rows=None, alg_k=None, gn_rpi=0, ln_rows=None, ln_in=None, ln_colsum=None, ln_eps=1e-5):
        """``gn_rpi`` > 0: the epilogue also writes GroupNorm partial statistics of ``out`` (rows of M space per image);
        ``ln_rows``: ... LayerNorm row partials of ``out``; ``ln_in`` / ``ln_colsum``: LayerNorm(A) folded in."""
        M = A.shape[0] if M is None else M
        gn_part = None
        if gn_rpi and FUSED_GN_STATS and not out_fp32 and W.shape[0] % 32 == 0:
            gn_part = self.buf(f"gnp_{out.data_ptr()}_{W.shape[0]}", ops.gn_partial_floats(self.N, gn_rpi, W.shape[0]), 1, F32)
            self.gn_src[(out.data_ptr(), W.shape[0])] = (gn_part, gn_rpi)
        else:
            self.gn_src.pop((out.data_ptr(), W.shape[0]), None)  # a stale record must not describe new contents
        g = ops.make_gemm(A_ptr=A.data_ptr(), lda=A.stride(0), a_rows=A.shape[0], W_ptr=W.data_ptr(),
                          ldw=W.stride(0), M=M, N=W.shape[0], Kc=A.shape[1] if Kc is None else Kc,
                          out_ptr=out.data_ptr(), ldo=out.stride(0), ntaps=ntaps, tap_wp=tap_wp,
                          bias_ptr=None if bias is None else bias.data_ptr(),
                          rowbias_ptr=None if rowbias is None else rowbias.data_ptr(), rowbias_rpi=rowbias_rpi,
                          residual_ptr=None if residual is None else residual.data_ptr(),
                          ldr=0 if residual is None else residual.stride(0), out_fp32=out_fp32, geglu=geglu,
                          out_map=out_map, H=H, Wd=Wd, mask_border=mask_border, Hp=Hp, Wp=Wp, bn=bn,
                          reverse=self._dir(A, out),
                          gn_partial_ptr=None if gn_part is None else gn_part.data_ptr(), gn_rpi=gn_rpi,
                          ln_rowstats_ptr=None if ln_rows is None else ln_rows.data_ptr(),
                          ln_in_ptr=None if ln_in is None else ln_in.data_ptr(),
                          ln_in_slots=0 if ln_in is None else ln_in.shape[0], ln_eps=ln_eps,
                          ln_colsum_ptr=None if ln_colsum is None else ln_colsum.data_ptr())
        self.keep += [g, A, W, out, bias, rowbias, residual, ln_rows, ln_in, ln_colsum]
        m_alg = M if alg_rows is None else alg_rows
        kk = (A.shape[1] if Kc is None else Kc) if alg_k is None else alg_k
        self.add(self.lib.seva_gemm_bf16, C.byref(g), flops=2.0 * m_alg * W.shape[0] * kk * ntaps,
                 tag=("conv3x3" if ntaps == 9 else ("upconv" if ntaps == 4 else ("geglu" if geglu else "linear")))
                 + f" M{M} N{W.shape[0]} K{int(kk * ntaps)}")

    def conv3x3(self, x, l, W9, bias, out, *, rowbias=None, residual=None, out_fp32=False, alg_k=None, stats=True):
        H, Wd = self.HW[l]
        self.gemm(x, W9, out, bias=bias, ntaps=9, tap_wp=Wd + 2, rowbias=rowbias, rowbias_rpi=(H + 2) * (Wd + 2),
                  residual=residual, out_fp32=out_fp32, mask_border=True, Hp=H + 2, Wp=Wd + 2,
                  alg_rows=self.rows_c(l), alg_k=alg_k, gn_rpi=(H + 2) * (Wd + 2) if stats and not out_fp32 else 0)

    def gn(self, x, l, C_, gamma, beta, eps, out, *, in_padded=True, out_padded=True, silu=False, dense=None,
           dense_w=None, dense_b=None):
        """GroupNorm: stats launch (+ memset inside) and apply launch."""
        H, Wd = self.HW[l]
        rpi = (H + 2) * (Wd + 2) if in_padded else H * Wd
        act = 2.0 * self.N * H * Wd * C_
        srcs = self._gn_sources(x, C_)
        if srcs is not None:
            # statistics came out of the producing GEMM epilogues: fold their partials (no pass over x)
            gst = self.buf(f"gn_gstats_{C_}", self.N * 32 * 2, 1, F32)
            (p0, c0, r0), (p1, c1, r1) = srcs[0], (srcs[1] if len(srcs) > 1 else (None, 0, 0))
            self.add(self.lib.seva_groupnorm_finalize, p0.data_ptr(), c0, r0, None if p1 is None else p1.data_ptr(), c1, r1,
                     self.N, 32, H * Wd, eps, gst.data_ptr(), nbytes=act / 8.0, tag=f"gn_finalize HW{H * Wd} C{C_}")
            pa = ops.make_gn_apply(x_ptr=x.data_ptr(), ldx=x.stride(0), y_ptr=out.data_ptr(), ldy=out.stride(0),
                                   sums_ptr=None, gamma_ptr=gamma.data_ptr(), beta_ptr=beta.data_ptr(),
                                   n_img=self.N, H=H, Wd=Wd, Cc=C_, groups=32, eps=eps, in_padded=in_padded,
                                   out_padded=out_padded, silu=silu,
                                   dense_ptr=None if dense is None else dense.data_ptr(),
                                   dense_w_ptr=None if dense_w is None else dense_w.data_ptr(),
                                   dense_b_ptr=None if dense_b is None else dense_b.data_ptr(),
                                   reverse=self._dir(x, out), gstats_ptr=gst.data_ptr())
            self.keep += [pa, x, out, gamma, beta, dense, dense_w, dense_b, gst]
            self.add(self.lib.seva_groupnorm_apply, C.byref(pa), nbytes=2.0 * act,
                     tag=f"gn_apply HW{H * Wd} C{C_} d{int(dense is not None)}")
            return
        # one buffer per (rows, channels): its ticket words sit at a shape-dependent offset and must stay zero between
        # launches, so shapes never share a buffer (zero-initialised by ``buf``, self-resetting afterwards)
        sums = self.buf(f"gn_sums_{rpi}_{C_}", ops.gn_stats_floats(self.N, rpi, C_), 1, F32)
        self.add(self.lib.seva_groupnorm_stats, x.data_ptr(), x.stride(0), self.N, rpi, C_, sums.data_ptr(),
                 int(self._dir(x)), nbytes=act, tag=f"gn_stats HW{H * Wd} C{C_}")
        pa = ops.make_gn_apply(x_ptr=x.data_ptr(), ldx=x.stride(0), y_ptr=out.data_ptr(), ldy=out.stride(0),
                               sums_ptr=sums.data_ptr(), gamma_ptr=gamma.data_ptr(), beta_ptr=beta.data_ptr(),
                               n_img=self.N, H=H, Wd=Wd, Cc=C_, groups=32, eps=eps, in_padded=in_padded,
                               out_padded=out_padded, silu=silu,
                               dense_ptr=None if dense is None else dense.data_ptr(),
                               dense_w_ptr=None if dense_w is None else dense_w.data_ptr(),
                               dense_b_ptr=None if dense_b is None else dense_b.data_ptr(),
                               reverse=self._dir(x, out))
        self.keep += [pa, x, out, gamma, beta, dense, dense_w, dense_b]
        self.add(self.lib.seva_groupnorm_apply, C.byref(pa), nbytes=2.0 * act, tag=f"gn_apply HW{H * Wd} C{C_} d{int(dense is not None)}")

    def _gn_sources(self, x, C_):
        """[(partial, channels, rows per image)] of the one or two GEMM outputs (side by side along the channel axis)
        that make up ``x``, or None when some part of it was not produced with statistics."""
        got, ptr, left = [], x.data_ptr(), C_
        while left > 0:
            hit = next(((k, v) for k, v in self.gn_src.items() if k[0] == ptr and k[1] <= left), None)
            if hit is None or len(got) == 2:
                return None
            (_, ncols), (part, rpi) = hit
            got.append((part, ncols, rpi))
            ptr += 2 * ncols
            left -= ncols
        return got

    def ln(self, x, nrm, out):
        self.add(self.lib.seva_layernorm, x.data_ptr(), x.stride(0), out.data_ptr(), out.stride(0),
                 nrm["g"].data_ptr(), nrm["b"].data_ptr(), x.shape[0], x.shape[1], nrm["eps"], int(self._dir(x, out)),
                 nbytes=4.0 * x.shape[0] * x.shape[1], tag=f"layernorm M{x.shape[0]} C{x.shape[1]}")
        self.keep += [x, out]

    def gemv(self, x, W, bias, out, *, silu_in=False, silu_out=False, group=None):
        """Small-M Linear.  With ``group`` the problem joins a grouped launch that is emitted once, right after
        the time embedding (all grouped problems depend only on the step's (t, context), not on activations)."""
        M, K = x.shape
        self.keep += [x, W, bias, out]
        if group is None:
            self.add(self.lib.seva_gemv_rows, x.data_ptr(), W.data_ptr(), None if bias is None else bias.data_ptr(),
                     out.data_ptr(), M, W.shape[0], K, int(silu_in), int(silu_out), tag="gemv")
            return
        pr = _lib.SevaGemvProblem()
        pr.x, pr.W, pr.bias, pr.out = x.data_ptr(), W.data_ptr(), None if bias is None else bias.data_ptr(), out.data_ptr()
        pr.M, pr.N, pr.K, pr.silu_in, pr.silu_out = M, W.shape[0], K, int(silu_in), int(silu_out)
        self.gemv_groups.setdefault(group, []).append(pr)

    def _emit_gemv_groups(self, pos: int) -> None:
        """Insert one grouped launch per dependency level at plan position ``pos``."""
        for name in ("emb", "cross_v", "cross_o"):
            probs = self.gemv_groups.get(name)
            if not probs:
                continue
            arr = (_lib.SevaGemvProblem * len(probs))(*probs)
            table = torch.frombuffer(bytearray(bytes(arr)), dtype=torch.uint8).to(self.dev)
            self.keep.append(table)
            args = (table.data_ptr(), len(probs), max(p.M for p in probs), max(p.N for p in probs),
                    max(p.K for p in probs))
            self.calls.insert(pos, (self.lib.seva_gemv_grouped, args))
            self.meta.insert(pos, ("seva_gemv_grouped", name, 0.0, 0.0))
            self.n_launch += 1
            self._invariant = {i + 1 if i >= pos else i for i in self._invariant}
            self._emb_calls = {i + 1 if i >= pos else i for i in self._emb_calls}
            if name != "emb":  # the cross-attention vectors depend on the context only, not on the step
                self._invariant.add(pos)
            else:
                self._emb_calls.add(pos)
            pos += 1
        if self._emb_outs:
            # the table path's only launch: rows of the per-index table -> the ResBlocks' row-bias buffers
            desc = torch.zeros(len(self._emb_outs), 3, dtype=torch.int64)
            col = 0
            for i, (buf, cout) in enumerate(self._emb_outs):
                desc[i, 0], desc[i, 1], desc[i, 2] = col, cout, buf.data_ptr()
                col += cout
            self._emb_desc = desc.to(self.dev)
            self._emb_cols = col
            self._emb_gather_idx = pos
            self.calls.insert(pos, (_no_launch, ()))   # replaced when the table exists; a no-op until then
            self.meta.insert(pos, ("seva_emb_gather", "emb table", 0.0, 0.0))
            self._invariant = {i + 1 if i >= pos else i for i in self._invariant}
            self._emb_calls = {i + 1 if i >= pos else i for i in self._emb_calls}

    # ---- network -------------------------------------------------------------------------------
    def _build(self, model: Seva):
        p = model.params
        N, lib = self.N, self.lib
        mc, ted = p.model_channels, p.model_channels * 4
        # time embedding (layers.py:11-32, model.py:47-52,185-186)
        temb = self.buf("temb", N, mc, F32)
        e1 = self.buf("temb1", N, ted, F32)
        self.emb = self.buf("emb", N, ted, F32)
        te = self.P["time_embed"]
        self._emb_calls.update(range(len(self.calls), len(self.calls) + 3))
        self.add(lib.seva_timestep_embedding, self.t_in.data_ptr(), temb.data_ptr(), N, mc, 10000.0)
        self.gemv(temb, te["w0"], te["b0"], e1, silu_out=True)
        # ``emb`` is consumed only through ``emb_layers = Sequential(SiLU, Linear)`` (layers.py:102-105): the SiLU is
        # applied once here instead of once per ResBlock and column tile
        self.gemv(e1, te["w2"], te["b2"], self.emb, silu_out=True)
        # Plucker map resized once per level (layers.py:126-128), channels-last fp32
        self.dense = []
        for l, (H, Wd) in enumerate(self.HW):
            d = self.buf(f"dense{l}", N, H * Wd * p.dense_in_channels, F32)
            self._invariant.add(len(self.calls))  # depends on the Plucker map only
            self.add(lib.seva_resize_bilinear_ac, self.dense_in.data_ptr(), d.data_ptr(), N, p.dense_in_channels,
                     self.h, self.w, H, Wd)
            self.dense.append(d)
        group_pos = len(self.calls)

        # ---- static schedule of where every block output lives (virtual skip concat) -----------
        in_blocks = list(model.input_blocks)
        out_blocks = list(model.output_blocks)
        in_level, in_ch = [], []
        lvl, ch = 0, mc
        for i, blk in enumerate(in_blocks):
            for layer in blk:
                if isinstance(layer, ResBlock):
                    ch = layer.out_channels
                elif isinstance(layer, Downsample):
                    lvl += 1
                    ch = layer.out_channels
            in_level.append(lvl)
            in_ch.append(ch)
        h_dest, skip_dest, cats = [], [None] * len(in_blocks), []
        ch_h, lvl_o = ch, lvl
        for k, blk in enumerate(out_blocks):
            i = len(in_blocks) - 1 - k
            assert in_level[i] == lvl_o, "skip / decoder resolution mismatch"
            cat = self.buf(f"cat{k}", self.rows_p(lvl_o), ch_h + in_ch[i])
            cats.append(cat)
            h_dest.append(cat[:, :ch_h])
            skip_dest[i] = cat[:, ch_h:]
            for layer in blk:
                if isinstance(layer, ResBlock):
                    ch_h = layer.out_channels
                elif isinstance(layer, Upsample):
                    lvl_o -= 1
        final_h = self.buf("final_h", self.rows_p(0), ch_h)

        # ---- encoder ----------------------------------------------------------------------------
        ci = self.P["conv_in"]
        self.conv3x3(self.net_in, 0, ci["w"], ci["b"], skip_dest[0], alg_k=p.in_channels)
        x, lvl = skip_dest[0], 0
        for i in range(1, len(in_blocks)):
            x, lvl = self._block(in_blocks[i], x, lvl, skip_dest[i], f"in{i}")
        # ---- middle -----------------------------------------------------------------------------
        x, lvl = self._block(model.middle_block, x, lvl, h_dest[0], "mid")
        # ---- decoder ----------------------------------------------------------------------------
        for k, blk in enumerate(out_blocks):
            dest = h_dest[k + 1] if k + 1 < len(out_blocks) else final_h
            x, lvl = self._block(blk, cats[k], lvl, dest, f"out{k}")
        # ---- head (model.py:170-174) ------------------------------------------------------------
        po = self.P["out"]
        hn = self.buf("final_n", self.rows_p(0), ch_h)
        self.gn(final_h, 0, ch_h, po["g"], po["b"], po["eps"], hn, silu=True)
        self.conv3x3(hn, 0, po["w"], po["cb"], self.eps, out_fp32=True)
        self._emit_gemv_groups(group_pos)

    def _block(self, blk, x, lvl, dest, tag):
        layers = list(blk)
        for j, layer in enumerate(layers):
            last = j == len(layers) - 1
            if isinstance(layer, ResBlock):
                out = dest if last else self.buf(f"res_out_{lvl}", self.rows_p(lvl), layer.out_channels)
                self._resblock(layer, x, lvl, out)
            elif isinstance(layer, MultiviewTransformer):
                out = dest if last else self.buf(f"mvt_out_{lvl}", self.rows_p(lvl), layer.in_channels)
                self._mvt(layer, x, lvl, out)
            elif isinstance(layer, Downsample):
                out = dest
                self._down(layer, x, lvl, out)
                lvl += 1
            elif isinstance(layer, Upsample):
                out = dest
                self._up(layer, x, lvl, out)
                lvl -= 1
            else:
                raise SevaB200Error(f"unexpected layer {type(layer).__name__} in {tag}")
            x = out
        return x, lvl

    # ResBlock.forward, layers.py:120-139
    def _resblock(self, m: ResBlock, x, l, out):
        w = self.P[id(m)]
        cin, cout = m.channels, m.out_channels
        rp = self.rows_p(l)
        h1 = self.buf(f"rb_h1_{l}", rp, cin)
        self.gn(x, l, cin, w["g1"], w["b1"], w["eps1"], h1, silu=True, dense=self.dense[l], dense_w=w["wd"],
                dense_b=w["bd"])
        emb_out = self.buf(f"rb_emb_{id(m)}", self.N, cout, F32)
        self.gemv(self.emb, w["we"], w["be"], emb_out, group="emb")
        self._emb_outs.append((emb_out, cout))
        h2 = self.buf(f"rb_h2_{l}", rp, cout)
        self.conv3x3(h1, l, w["w1"], w["cb1"], h2, rowbias=emb_out)
        h3 = self.buf(f"rb_h3_{l}", rp, cout)
        self.gn(h2, l, cout, w["g2"], w["b2"], w["eps2"], h3, silu=True)
        if "ws" in w:
            H, Wd = self.HW[l]
            skip = self.buf(f"rb_skip_{l}", rp, cout)
            self.gemm(x, w["ws"], skip, bias=w["bs"], mask_border=True, Hp=H + 2, Wp=Wd + 2,
                      alg_rows=self.rows_c(l))
        else:
            skip = x
        self.conv3x3(h3, l, w["w2"], w["cb2"], out, residual=skip)

    # Downsample.forward, layers.py:56-58: stride-2 conv = gather + GEMM, written into the next level's border image
    def _down(self, m: Downsample, x, l, out):
        w = self.P[id(m)]
        H, Wd = self.HW[l]
        Ho, Wo = self.HW[l + 1]
        col = self.buf(f"im2col_{l}", self.N * Ho * Wo, 9 * m.channels)
        self.add(self.lib.seva_im2col3x3_s2, x.data_ptr(), x.stride(0), col.data_ptr(), self.N, H, Wd, m.channels)
        self._touch_forward(x, col)
        self.keep += [x]
        self.gemm(col, w["w"], out, bias=w["b"], out_map=1, H=Ho, Wd=Wo, gn_rpi=Ho * Wo)

    # Upsample.forward, layers.py:42-46
    def _up(self, m: Upsample, x, l, out):
        """Nearest-2x upsampling + 3x3 conv as four 2x2-tap phase convolutions on the SOURCE image (SevaGemm.out_map = 2..5):
        16/36 of the multiply-adds of the upsample-then-convolve form and no upsampled copy.  ``out`` keeps the zero
        border it was allocated with (no launch of the plan writes non-zero border rows)."""
        w = self.P[id(m)]
        H, Wd = self.HW[l]
        cin = m.channels
        for ph in range(4):
            self.gemm(x, w["wph"][ph], out, bias=w["b"], ntaps=4, tap_wp=Wd + 2, out_map=2 + ph, H=H, Wd=Wd, Kc=cin,
                      alg_rows=self.N * H * Wd, alg_k=9.0 * cin / 4.0)

    # cross-attention with a single context token is a per-image vector: softmax over one key is 1, so
    # attn2(x, ctx) = to_out(to_v(ctx)) for every query (transformer.py:59-74 with Lk = 1).
    def _cross_vec(self, wa, ctx_rows, C_, tag):
        R = ctx_rows.shape[0]
        v = self.buf(f"xv_{tag}", R, C_, F32)
        o = self.buf(f"xo_{tag}", R, C_, F32)
        self.gemv(ctx_rows, wa["wv1"], None, v, group="cross_v")
        self.gemv(v, wa["wo1"], wa["bo"], o, group="cross_o")
        return o

    def _rows(self, l, C_):
        """LayerNorm row-partial buffer of level ``l`` ([slots, rows, 2] fp32): written by the epilogue of the GEMM
        that produces a LayerNorm input and read by the GEMM the LayerNorm is folded into, back to back."""
        return self.buf(f"tk_rows_{l}", ops.ln_slots(C_) * self.rows_c(l) * 2, 1, F32).view(ops.ln_slots(C_), self.rows_c(l), 2)

    def _ff(self, x, nrm, ff, residual, out, l, C_in, tag, x_rows=None, out_rows=None):
        """FeedForward(LayerNorm(x)) + residual.  ``x_rows``: row statistics of x from its producer (the LayerNorm is
        then folded into the GEGLU projection); ``out_rows``: have the down projection write row statistics of out."""
        rc = x.shape[0]
        inner = ff["w1"].shape[0] // 2
        g = self.buf(f"tk_g_{l}", rc, inner)
        if x_rows is not None:
            self.gemm(x, ff["w1"], g, bias=ff["b1"], geglu=True, bn=ff["bn1"], ln_in=x_rows, ln_colsum=ff["cs1"],
                      ln_eps=ff["eps"])
        else:
            a = self.buf(f"tk_ln_{l}", rc, C_in)
            self.ln(x, nrm, a)
            self.gemm(a, ff["w1"], g, bias=ff["b1"], geglu=True, bn=ff["bn1"])
        self.gemm(g, ff["w2"], out, bias=ff["b2"], residual=residual, ln_rows=out_rows)

    # MultiviewTransformer.forward, transformer.py:215-247
    def _mvt(self, m: MultiviewTransformer, x, l, out):
        w = self.P[id(m)]
        C_, heads = m.in_channels, m.n_heads
        # heads narrower than 64 run zero-padded to 64 (packed weights carry the padding; the padded score / output
        # dimensions contribute exact zeros): Ci is the padded inner width, the softmax scale stays 1 / sqrt(d_head)
        Ci, sc = heads * 64, 1.0 / math.sqrt(m.d_head)
        self._attn_scale = sc
        H, Wd = self.HW[l]
        S = H * Wd
        rc = self.rows_c(l)
        is3d = m.name in self.unflatten
        T, B = self.T, self.B
        ctx2 = self.ctx_in.reshape(self.N, self.L * self.ctx_dim)
        ctx0 = self.ctx0_in.reshape(B, self.L * self.ctx_dim)

        xn = self.buf(f"tk_xn_{l}", rc, C_)
        self.gn(x, l, C_, w["gn"]["g"], w["gn"]["b"], w["gn"]["eps"], xn, out_padded=False)
        x0 = self.buf(f"tk_x0_{l}", rc, C_)
        # LayerNorm -> Linear pairs: the producer of the LayerNorm input writes per-row (sum, sum of squares) partials
        # in its epilogue and the consuming projection applies mean / rstd in ITS epilogue (gamma / beta folded into
        # the packed weights), so no LayerNorm pass touches HBM.  ``rows`` is None when folding is off for this width.
        fold = w["ln_fold"]
        rows = self._rows(l, C_) if fold else None
        self.gemm(xn, w["win"], x0, bias=w["bin"], ln_rows=rows)

        # transformer_depth (block, mix_block) pairs (transformer.py:234-241); the buffers are reused pair after pair
        for d, pw in enumerate(w["pairs"]):
            more = d + 1 < len(w["pairs"])
            # ---- spatial block (TransformerBlock.forward, transformer.py:106-110)
            a = self.buf(f"tk_ln_{l}", rc, C_)
            if not fold:
                self.ln(x0, pw["s_n1"], a)
            qkv_in = dict(A=x0, ln=dict(ln_in=rows, ln_eps=pw["s_attn1"]["eps"])) if fold else dict(A=a, ln={})
            qkv = self.buf(f"tk_qkv_{l}", rc, 3 * Ci)
            o = self.buf(f"tk_o_{l}", rc, Ci)
            bat, Lseq = (B, T * S) if is3d else (self.N, S)
            sh = self.shard
            if is3d and sh is not None:
                # frame sequence-parallel 3D attention: local queries against the keys / values of all Tk frames
                ring = os.environ.get("SEVA_RING", "ring")   # ring | allgather (both overlapped) | off (blocking gather)
                Lk = self.Tk * S
                if ring != "off":
                    box = {}

                    def after_kv(q_, kvl_):
                        box["ring"] = _RingAttention(self.lib, sh, q_, kvl_, o, bat, heads, S, Ci, mode=ring, scale=sc)
                        self.add(_RingStart(box["ring"]), tag="start")

                    q, kvl, _ = self._split_qkv(qkv_in, pw["s_attn1"], l, Ci, S, gather=False, after_kv=after_kv)
                    self.add(box["ring"],
                             flops=4.0 * bat * heads * Lseq * Lk * 64, nbytes=2.0 * bat * Lk * 2 * Ci,
                             tag=f"attn3d+exchange B{bat} H{heads} Lq{Lseq} Lk{Lk} {ring}")
                    self._touch_forward(q, o)
                else:
                    q, kvl, kvf = self._split_qkv(qkv_in, pw["s_attn1"], l, Ci, S)
                    self.add(self.lib.seva_attn_fwd, q.data_ptr(), kvf.data_ptr(), kvf.data_ptr() + 2 * Ci, o.data_ptr(), bat,
                             heads, Lseq, Lk, Ci, 2 * Ci, 2 * Ci, o.stride(0), Lseq * Ci, Lk * 2 * Ci, Lk * 2 * Ci,
                             Lseq * o.stride(0), sc, flops=4.0 * bat * heads * Lseq * Lk * 64,
                             tag=f"attn3d B{bat} H{heads} Lq{Lseq} Lk{Lk}")
                    self._touch_forward(q, kvf, o)
            else:
                self._qkv(qkv_in, pw["s_attn1"], slice(None), qkv)
                ld = qkv.stride(0)
                base = qkv.data_ptr()
                self.add(self.lib.seva_attn_fwd, base, base + 2 * Ci, base + 4 * Ci, o.data_ptr(), bat, heads, Lseq, Lseq,
                         ld, ld, ld, o.stride(0), Lseq * ld, Lseq * ld, Lseq * ld, Lseq * o.stride(0), sc,
                         flops=4.0 * bat * heads * Lseq * Lseq * 64,
                         tag=("attn3d" if is3d else "attn2d") + f" B{bat} H{heads} L{Lseq}")
                self._touch_forward(qkv, o)
            x1 = self.buf(f"tk_x1_{l}", rc, C_)
            if self.L == 1:
                xr = self._cross_vec(pw["s_attn2"], ctx0 if is3d else ctx2, C_, f"s{id(m)}_{d}")
                self.gemm(o, pw["s_attn1"]["wo"], x1, bias=pw["s_attn1"]["bo"], residual=x0, rowbias=xr, rowbias_rpi=Lseq,
                          ln_rows=rows)
            else:
                x1a = self.buf(f"tk_x1a_{l}", rc, C_)
                self.gemm(o, pw["s_attn1"]["wo"], x1a, bias=pw["s_attn1"]["bo"], residual=x0)
                self._cross_general(pw["s_attn2"], pw["s_n2"], x1a, x1, l, Ci, heads, bat, Lseq,
                                    self.ctx0_in if is3d else self.ctx_in, out_rows=rows)
            x2 = self.buf(f"tk_x2_{l}", rc, C_)
            self._ff(x1, pw["s_n3"], pw["s_ff"], x1, x2, l, C_, "s", x_rows=rows, out_rows=rows)

            # ---- time-mix block (TransformerBlockTimeMix.forward, transformer.py:145-155), tokens stay (b t) s
            y1 = self.buf(f"tk_y1_{l}", rc, C_)
            self._ff(x2, pw["t_nin"], pw["t_ffin"], x2, y1, l, C_, "tin", x_rows=rows, out_rows=rows)
            if not fold:
                self.ln(y1, pw["t_n1"], a)
            qkv_in = dict(A=y1, ln=dict(ln_in=rows, ln_eps=pw["t_attn1"]["eps"])) if fold else dict(A=a, ln={})
            if sh is not None and self.tm_a2a and S >= sh.size:
                sa, sb = sh.space_parts(S)[sh.rank]
                S_me = sb - sa
                self._qkv(qkv_in, pw["t_attn1"], slice(None), qkv)
                qkv_sp = self.buf(f"tk_qkv_sp_{l}", B * self.Tk * S_me, 3 * Ci)
                o_sp = self.buf(f"tk_o_sp_{l}", B * self.Tk * S_me, Ci)
                self.add(_SpaceExchange(sh, True, qkv, qkv_sp, B, S), nbytes=2.0 * (qkv.numel() + qkv_sp.numel()),
                         tag=f"to_space rows{rc} C{3 * C_}")
                self.add(self.lib.seva_timemix_attn, qkv_sp.data_ptr(), o_sp.data_ptr(), B, self.Tk, S_me, heads,
                         qkv_sp.stride(0), o_sp.stride(0), sc, flops=4.0 * B * S_me * heads * self.Tk * self.Tk * 64,
                         nbytes=2.0 * 4 * qkv_sp.shape[0] * C_, tag=f"timemix BS{B * S_me} H{heads} T{self.Tk}")
                self.add(_SpaceExchange(sh, False, o_sp, o, B, S), nbytes=2.0 * (o_sp.numel() + o.numel()),
                         tag=f"to_frames rows{rc} C{C_}")
                self._touch_forward(qkv, o)
            elif sh is not None:
                q, kvl, kvf = self._split_qkv(qkv_in, pw["t_attn1"], l, Ci, S)
                self.add(self.lib.seva_timemix_attn_kv, q.data_ptr(), kvf.data_ptr(), kvf.data_ptr() + 2 * Ci,
                         o.data_ptr(), B, T, self.Tk, S, heads, Ci, 2 * Ci, o.stride(0), sc,
                         flops=4.0 * B * S * heads * T * self.Tk * 64, nbytes=2.0 * (2 * rc + 2 * B * self.Tk * S) * C_,
                         tag=f"timemix BS{B * S} H{heads} Tq{T} Tk{self.Tk}")
                self._touch_forward(q, kvf, o)
            else:
                self._qkv(qkv_in, pw["t_attn1"], slice(None), qkv)
                self.add(self.lib.seva_timemix_attn, qkv.data_ptr(), o.data_ptr(), B, T, S, heads, qkv.stride(0),
                         o.stride(0), sc, flops=4.0 * B * S * heads * T * T * 64, nbytes=2.0 * 4 * rc * C_,
                         tag=f"timemix BS{B * S} H{heads}")
                self._touch_forward(qkv, o)
            y2 = self.buf(f"tk_y2_{l}", rc, C_)
            if self.L == 1:
                xr = self._cross_vec(pw["t_attn2"], ctx0, C_, f"t{id(m)}_{d}")
                self.gemm(o, pw["t_attn1"]["wo"], y2, bias=pw["t_attn1"]["bo"], residual=y1, rowbias=xr,
                          rowbias_rpi=T * S, ln_rows=rows)
            else:
                # time_context = context[::T] repeated over the positions (transformer.py:222-226): every token of batch
                # element b attends to the same L context tokens, and cross-attention is per query token, so the (b t) s
                # token order needs no transpose: batch = B, T * S queries each
                y2a = self.buf(f"tk_x1a_{l}", rc, C_)
                self.gemm(o, pw["t_attn1"]["wo"], y2a, bias=pw["t_attn1"]["bo"], residual=y1)
                self._cross_general(pw["t_attn2"], pw["t_n2"], y2a, y2, l, Ci, heads, B, T * S, self.ctx0_in, out_rows=rows)
            x3 = self.buf(f"tk_x3_{l}", rc, C_)
            # x = x_spatial + ff(norm3(y2)): SkipConnect (transformer.py:158-165) folded into the residual
            self._ff(y2, pw["t_n3"], pw["t_ff"], x2, x3, l, C_, "t", x_rows=rows, out_rows=rows if more else None)
            x0 = x3   # input of the next (block, mix_block) pair
        # proj_out + input residual, written back to the zero-bordered layout (transformer.py:244-247); its epilogue
        # also leaves the GroupNorm partials of the block output for the next ResBlock
        self.gemm(x3, w["wout"], out, bias=w["bout"], residual=x, out_map=1, H=H, Wd=Wd, gn_rpi=S)

    def _qkv(self, src, wa, rows_sel, out):
        """Rows ``rows_sel`` of the fused q|k|v projection of ``src['A']`` (the LayerNorm output, or the LayerNorm
        INPUT with its row statistics when the LayerNorm is folded into this projection)."""
        if src["ln"]:
            self.gemm(src["A"], wa["wqkv"][rows_sel], out, bias=wa["bqkv"][rows_sel], ln_colsum=wa["csqkv"][rows_sel],
                      **src["ln"])
        else:
            self.gemm(src["A"], wa["wqkv"][rows_sel], out)

    def _split_qkv(self, src, wa, l, C_, S, gather=True, after_kv=None):
        """Frame-sharded self-attention input: q for the local frames, [k | v] gathered over the frame group
        (``gather=False``: only the local [k | v]; the caller exchanges it itself).  K/V are projected first so that
        an overlapped exchange can start while q is still being projected."""
        rc = src["A"].shape[0]
        q = self.buf(f"tk_q_{l}", rc, C_)
        kvl = self.buf(f"tk_kvl_{l}", rc, 2 * C_)
        self._qkv(src, wa, slice(C_, None), kvl)
        if after_kv is not None:
            after_kv(q, kvl)
        self._qkv(src, wa, slice(0, C_), q)
        if not gather:
            return q, kvl, None
        kvf = self.buf(f"tk_kvf_{l}", self.B * self.Tk * S, 2 * C_)
        self.add(_FrameGather(self.shard, kvf, kvl, self.B, S), nbytes=2.0 * kvf.numel(), tag=f"kv rows{kvf.shape[0]} C{2 * C_}")
        return q, kvl, kvf

    def _cross_general(self, wa, nrm, xin, xout, l, C_, heads, bat, Lseq, ctx, out_rows=None):
        """attn2 with Lk > 1 context tokens: LN -> q, k/v from context -> flash attention -> to_out + residual."""
        rc = xin.shape[0]
        sc = self._attn_scale
        a = self.buf(f"tk_ln_{l}", rc, xin.shape[1])
        self.ln(xin, nrm, a)
        q = self.buf(f"tk_q_{l}", rc, C_)
        self.gemm(a, wa["wq"], q)
        R = ctx.shape[0]
        cb = self.buf(f"ctx_bf16_{R}", R * self.L, self.ctx_dim)
        self.add(self.lib.seva_cast_f32_bf16, ctx.data_ptr(), cb.data_ptr(), ctx.numel())
        k = self.buf(f"ctx_k_{R}_{C_}", R * self.L, C_)
        v = self.buf(f"ctx_v_{R}_{C_}", R * self.L, C_)
        self.gemm(cb, wa["wk"], k)
        self.gemm(cb, wa["wv"], v)
        o = self.buf(f"tk_o2_{l}", rc, C_)
        assert R == bat
        self.add(self.lib.seva_attn_fwd, q.data_ptr(), k.data_ptr(), v.data_ptr(), o.data_ptr(), bat, heads, Lseq,
                 self.L, C_, C_, C_, C_, Lseq * C_, self.L * C_, self.L * C_, Lseq * C_, sc,
                 flops=4.0 * bat * heads * Lseq * self.L * 64, tag="attn_cross")
        self._touch_forward(q, o)
        self.gemm(o, wa["wo"], xout, bias=wa["bo"], residual=xin, ln_rows=out_rows)

    # ---- execution ------------------------------------------------------------------------------
    def cond_current(self, epoch) -> bool:
        """True if the conditioning of ``epoch`` is already bound (the caller can skip preparing it)."""
        return epoch is not None and epoch == self._epoch

    def bind(self, t: torch.Tensor, ctx: torch.Tensor | None, dense: torch.Tensor | None,
             ctx0: torch.Tensor | None = None, epoch=None, t_range: int = 0) -> None:
        """Copy the per-call conditioning into the plan's static input buffers (device-to-device).  ``ctx0`` is
        ``context[::T]`` (transformer.py:222-229); a frame shard passes the full batch's, since frame 0 may live
        on another rank.  ``epoch`` (from the guider's ``prepare_inputs`` cache, one value per sampling loop and
        conditioning) marks ctx / dense as unchanged since the last call with the same value: they are not copied
        again and the launches that depend only on them are skipped by ``run``."""
        self.t_in.copy_(t.reshape(-1))
        # ``t_range`` > 0: the caller guarantees 0 <= t < t_range (the denoiser's sigma indices, sampling.py:126-131), so the
        # timestep-only launches may be served from the per-index table
        self._use_emb_table = bool(t_range) and self.emb_table_enabled and self._emb_gather_idx is not None
        if self._use_emb_table and (self._emb_table is None or self._emb_rows != t_range):
            self._build_emb_table(int(t_range))
        if self.cond_current(epoch):
            self._skip_invariant = True
            return
        self.ctx_in.copy_(ctx)
        self.ctx0_in.copy_(ctx[:: self.T] if ctx0 is None else ctx0)
        self.dense_in.copy_(dense)
        self._epoch = epoch
        self._skip_invariant = False

    @property
    def n_launch_step(self) -> int:
        """Launches of a step whose conditioning is already bound."""
        return self.n_launch - len(self._invariant) - (len(self._emb_calls) - 1 if self._use_emb_table else 0)

    def _build_emb_table(self, rows: int) -> None:
        """N2: ``emb_layers(time_embed(timestep_embedding(t)))`` for every sigma index, computed ONCE per engine with the
        plan's own launches (N indices at a time, so every table row is bit-identical to what a step would compute), kept as
        fp32 [rows, sum of the ResBlocks' widths] (80 MB at the released size).  A step then gathers its rows
        (``seva_emb_gather``) instead of running the t-emb kernel, two GEMVs and the grouped GEMV."""
        stream = torch.cuda.current_stream().cuda_stream
        table = torch.empty(rows, self._emb_cols, device=self.dev, dtype=F32)
        saved = self.t_in.clone()
        order = sorted(self._emb_calls)
        for r0 in range(0, rows, self.N):
            n = min(self.N, rows - r0)
            self.t_in.copy_(torch.arange(r0, r0 + self.N, device=self.dev).clamp_(max=rows - 1))
            for i in order:
                fn, args = self.calls[i]
                rc = fn(*args, stream)
                if rc != 0:
                    raise SevaB200Error(f"{fn.__name__} failed (rc={rc}): {_lib.last_error()}")
            col = 0
            for buf, cout in self._emb_outs:
                table[r0:r0 + n, col:col + cout].copy_(buf[:n])
                col += cout
        self.t_in.copy_(saved)
        self._emb_table, self._emb_rows = table, rows
        self.calls[self._emb_gather_idx] = (self.lib.seva_emb_gather, (table.data_ptr(), self._emb_cols, rows,
                                                                       self.t_in.data_ptr(), self.N,
                                                                       self._emb_desc.data_ptr(), len(self._emb_outs)))
        self._graphs.clear()
        self._eager_runs.clear()

    def _launch_all(self, skip, stream) -> None:
        for i, (fn, args) in enumerate(self.calls):
            if i in skip:
                continue
            rc = fn(*args, stream)
            if rc != 0:
                raise SevaB200Error(f"{fn.__name__} failed (rc={rc}): {_lib.last_error()}")

    def _capture(self, skip):
        """Record the plan (minus ``skip``) into a CUDA graph.  Nothing is allocated while capturing (every workspace
        exists already) and the launches only enqueue kernels, so the capture itself does no work: the caller replays."""
        g = torch.cuda.CUDAGraph()
        # thread_local: NCCL's watchdog thread may call the runtime while a sharded plan is being captured
        with torch.cuda.graph(g, capture_error_mode="thread_local"):
            self._launch_all(skip, torch.cuda.current_stream().cuda_stream)
        return g

    def run(self, timing: list | None = None) -> None:
        """Launch the plan on the current stream.  With ``timing`` (a list), a CUDA-event pair is recorded around
        every launch and (index, start, end) is appended; read ``self.meta[index]`` for name / algorithmic work."""
        if torch.cuda.current_device() != self.dev.index:  # pointers, stream and smem opt-ins all belong to self.dev
            with torch.cuda.device(self.dev):
                return self.run(timing)
        stream = torch.cuda.current_stream().cuda_stream
        timing = self.timing_sink if timing is None else timing
        skip = set(self._invariant) if self._skip_invariant else set()
        if self._emb_gather_idx is not None:
            skip |= self._emb_calls if self._use_emb_table else {self._emb_gather_idx}
        if timing is None:
            if self.use_graph:
                variant = (bool(self._skip_invariant), self._use_emb_table)
                g = self._graphs.get(variant)
                if g is None and self._eager_runs.get(variant, 0) >= 1 and not torch.cuda.is_current_stream_capturing():
                    g = self._capture(skip)
                    self._graphs[variant] = g
                if g is not None:
                    g.replay()
                    return
                self._eager_runs[variant] = self._eager_runs.get(variant, 0) + 1
            self._launch_all(skip, stream)
            return
        for i, (fn, args) in enumerate(self.calls):
            if i in skip:
                continue
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            rc = fn(*args, stream)
            e1.record()
            if rc != 0:
                raise SevaB200Error(f"{fn.__name__} failed (rc={rc}): {_lib.last_error()}")
            timing.append((i, e0, e1))
