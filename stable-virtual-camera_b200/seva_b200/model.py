"""Seva network on B200 kernels behind the reference's ``seva.model`` surface.

Mirrors (reference paths): ``SevaParams`` seva/model.py:17-36, ``Seva`` :39-216, ``SGMWrapper`` :219-234, and the
module tree of seva/modules/layers.py + seva/modules/transformer.py so that the 1146 state-dict keys are
identical (``load_state_dict(strict=False, assign=True)`` on a meta-initialised module works, utils.py:50-53).

The ``nn.Module`` tree below only *holds parameters*.  ``forward`` runs an execution plan (engine.py) made of
calls into the C ABI (include/seva_b200.h): activations stay channels-last bf16 on the device, convolutions
are implicit GEMMs on tcgen05, attention is the TMEM flash kernel.  No PyTorch compute op is on the path and
there is no CPU fallback: CPU tensors raise.
"""
from __future__ import annotations

from dataclasses import dataclass, field

import torch
import torch.nn as nn

from . import ops
from ._lib import SevaB200Error


@dataclass
class SevaParams(object):
    in_channels: int = 11
    model_channels: int = 320
    out_channels: int = 4
    num_frames: int = 21
    num_res_blocks: int = 2
    attention_resolutions: list[int] = field(default_factory=lambda: [4, 2, 1])
    channel_mult: list[int] = field(default_factory=lambda: [1, 2, 4, 4])
    num_head_channels: int = 64
    transformer_depth: list[int] = field(default_factory=lambda: [1, 1, 1, 1])
    context_dim: int = 1024
    dense_in_channels: int = 6
    dropout: float = 0.0
    unflatten_names: list[str] = field(default_factory=lambda: ["middle_ds8", "output_ds4", "output_ds2"])

    def __post_init__(self):
        assert len(self.channel_mult) == len(self.transformer_depth)


class _ParamHolder(nn.Module):
    """Parameter container: compute happens in the engine, never through nn.Module.forward."""

    def forward(self, *a, **k):  # pragma: no cover
        raise SevaB200Error(f"{type(self).__name__} holds parameters only; call Seva.forward / SGMWrapper.forward")


class GroupNorm32(nn.GroupNorm):
    pass


class TimestepEmbedSequential(nn.Sequential):
    pass


class Upsample(_ParamHolder):
    def __init__(self, channels: int, out_channels: int | None = None):
        super().__init__()
        self.channels, self.out_channels = channels, out_channels or channels
        self.conv = nn.Conv2d(self.channels, self.out_channels, 3, 1, 1)


class Downsample(_ParamHolder):
    def __init__(self, channels: int, out_channels: int | None = None):
        super().__init__()
        self.channels, self.out_channels = channels, out_channels or channels
        self.op = nn.Conv2d(self.channels, self.out_channels, 3, 2, 1)


class ResBlock(_ParamHolder):
    def __init__(self, channels, emb_channels, out_channels, dense_in_channels, dropout):
        super().__init__()
        out_channels = out_channels or channels
        self.channels, self.out_channels = channels, out_channels
        self.in_layers = nn.Sequential(GroupNorm32(32, channels), nn.SiLU(), nn.Conv2d(channels, out_channels, 3, 1, 1))
        self.emb_layers = nn.Sequential(nn.SiLU(), nn.Linear(emb_channels, out_channels))
        self.dense_emb_layers = nn.Sequential(nn.Conv2d(dense_in_channels, 2 * channels, 1, 1, 0))
        self.out_layers = nn.Sequential(GroupNorm32(32, out_channels), nn.SiLU(), nn.Dropout(dropout),
                                        nn.Conv2d(out_channels, out_channels, 3, 1, 1))
        self.skip_connection = nn.Identity() if out_channels == channels else nn.Conv2d(channels, out_channels, 1, 1, 0)


class GEGLU(_ParamHolder):
    def __init__(self, dim_in, dim_out):
        super().__init__()
        self.proj = nn.Linear(dim_in, dim_out * 2)


class FeedForward(_ParamHolder):
    def __init__(self, dim, dim_out=None, mult=4, dropout=0.0):
        super().__init__()
        inner = int(dim * mult)
        self.net = nn.Sequential(GEGLU(dim, inner), nn.Dropout(dropout), nn.Linear(inner, dim_out or dim))


class Attention(_ParamHolder):
    def __init__(self, query_dim, context_dim=None, heads=8, dim_head=64, dropout=0.0):
        super().__init__()
        self.heads, self.dim_head = heads, dim_head
        inner = heads * dim_head
        context_dim = context_dim or query_dim
        self.to_q = nn.Linear(query_dim, inner, bias=False)
        self.to_k = nn.Linear(context_dim, inner, bias=False)
        self.to_v = nn.Linear(context_dim, inner, bias=False)
        self.to_out = nn.Sequential(nn.Linear(inner, query_dim), nn.Dropout(dropout))


class TransformerBlock(_ParamHolder):
    def __init__(self, dim, n_heads, d_head, context_dim, dropout=0.0):
        super().__init__()
        self.attn1 = Attention(dim, None, n_heads, d_head, dropout)
        self.ff = FeedForward(dim, dropout=dropout)
        self.attn2 = Attention(dim, context_dim, n_heads, d_head, dropout)
        self.norm1, self.norm2, self.norm3 = nn.LayerNorm(dim), nn.LayerNorm(dim), nn.LayerNorm(dim)


class TransformerBlockTimeMix(_ParamHolder):
    def __init__(self, dim, n_heads, d_head, context_dim, dropout=0.0):
        super().__init__()
        inner = n_heads * d_head
        self.norm_in = nn.LayerNorm(dim)
        self.ff_in = FeedForward(dim, dim_out=inner, dropout=dropout)
        self.attn1 = Attention(inner, None, n_heads, d_head, dropout)
        self.ff = FeedForward(inner, dim_out=dim, dropout=dropout)
        self.attn2 = Attention(inner, context_dim, n_heads, d_head, dropout)
        self.norm1, self.norm2, self.norm3 = nn.LayerNorm(inner), nn.LayerNorm(inner), nn.LayerNorm(inner)


class SkipConnect(_ParamHolder):
    pass


class MultiviewTransformer(_ParamHolder):
    def __init__(self, in_channels, n_heads, d_head, name, unflatten_names=(), depth=1, context_dim=1024,
                 dropout=0.0):
        super().__init__()
        self.in_channels, self.name, self.unflatten_names = in_channels, name, list(unflatten_names)
        self.n_heads, self.d_head = n_heads, d_head
        inner = n_heads * d_head
        self.norm = nn.GroupNorm(32, in_channels, eps=1e-6)
        self.proj_in = nn.Linear(in_channels, inner)
        self.transformer_blocks = nn.ModuleList(
            [TransformerBlock(inner, n_heads, d_head, context_dim, dropout) for _ in range(depth)])
        self.proj_out = nn.Linear(inner, in_channels)
        self.time_mixer = SkipConnect()
        self.time_mix_blocks = nn.ModuleList(
            [TransformerBlockTimeMix(inner, n_heads, d_head, context_dim, dropout) for _ in range(depth)])


class Seva(nn.Module):
    def __init__(self, params: SevaParams) -> None:
        super().__init__()
        p = params
        if not (8 <= p.num_head_channels <= 64 and p.num_head_channels % 8 == 0):
            # the tcgen05 attention and time-mix kernels are built for 64-wide heads; narrower heads run zero-padded to 64
            # (engine._pad_head_rows), wider ones have no kernel
            raise SevaB200Error(f"num_head_channels={p.num_head_channels}: seva_b200 supports head widths 8, 16, ..., 64 "
                                "(64 is the released Seva configuration); there is no generic fallback")
        if p.model_channels % 64 != 0:
            raise SevaB200Error(f"model_channels={p.model_channels} must be a multiple of 64 (head width / GEMM K tiles)")
        self.params = p
        self.model_channels, self.out_channels, self.num_head_channels = p.model_channels, p.out_channels, p.num_head_channels
        mc = p.model_channels
        ted = mc * 4
        self.time_embed = nn.Sequential(nn.Linear(mc, ted), nn.SiLU(), nn.Linear(ted, ted))

        def res(cin, cout):
            return ResBlock(cin, ted, cout, p.dense_in_channels, p.dropout)

        def mvt(ch, name, level):
            return MultiviewTransformer(ch, ch // p.num_head_channels, p.num_head_channels, name=name,
                                        depth=p.transformer_depth[level], context_dim=p.context_dim,
                                        unflatten_names=p.unflatten_names)

        self.input_blocks = nn.ModuleList([TimestepEmbedSequential(nn.Conv2d(p.in_channels, mc, 3, padding=1))])
        self._feature_size = mc
        skip_chans = [mc]
        ch, ds = mc, 1
        last = len(p.channel_mult) - 1
        for level, mult in enumerate(p.channel_mult):
            for _ in range(p.num_res_blocks):
                layers = [res(ch, mult * mc)]
                ch = mult * mc
                if ds in p.attention_resolutions:
                    layers.append(mvt(ch, f"input_ds{ds}", level))
                self.input_blocks.append(TimestepEmbedSequential(*layers))
                self._feature_size += ch
                skip_chans.append(ch)
            if level != last:
                ds *= 2
                self.input_blocks.append(TimestepEmbedSequential(Downsample(ch, out_channels=ch)))
                skip_chans.append(ch)
                self._feature_size += ch
        self.middle_block = TimestepEmbedSequential(res(ch, None), mvt(ch, f"middle_ds{ds}", last), res(ch, None))
        self._feature_size += ch
        self.output_blocks = nn.ModuleList([])
        for level in range(last, -1, -1):
            mult = p.channel_mult[level]
            for i in range(p.num_res_blocks + 1):
                layers = [res(ch + skip_chans.pop(), mc * mult)]
                ch = mc * mult
                if ds in p.attention_resolutions:
                    layers.append(mvt(ch, f"output_ds{ds}", level))
                if level and i == p.num_res_blocks:
                    ds //= 2
                    layers.append(Upsample(ch, ch))
                self.output_blocks.append(TimestepEmbedSequential(*layers))
                self._feature_size += ch
        self.out = nn.Sequential(GroupNorm32(32, ch), nn.SiLU(), nn.Conv2d(mc, p.out_channels, 3, padding=1))
        self._engines: dict = {}
        self._packed = None
        self._frame_group = None
        self.register_load_state_dict_post_hook(lambda m, k: m.invalidate())

    # ---- engine management --------------------------------------------------------------------
    def invalidate(self) -> None:
        """Drop packed weights / plans (called on .to(), load_state_dict, or after editing parameters)."""
        self._engines = {}
        self._packed = None

    def _apply(self, fn, *a, **k):
        self.invalidate()
        return super()._apply(fn, *a, **k)

    def set_frame_group(self, group) -> None:
        """Frame sequence-parallel execution: the ranks of ``group`` split the ``num_frames`` frames of
        every forward (``dist.partition_frames``); callers keep passing, and receive, full replicated batches.  Per-frame
        ops run on the owner of the frame; K/V are exchanged at the 3D and time-mix self-attention.  ``None`` disables."""
        self._frame_group = group
        self._engines = {}

    def engine(self, n: int, num_frames: int, h: int, w: int, ctx_len: int):
        """Plan for a forward of ``n`` images in sets of ``num_frames`` (global sizes; with a frame group the plan
        covers this rank's frames and ``engine.shard`` describes the split)."""
        from .engine import Engine, FrameShard, pack_weights

        dev = next(self.parameters()).device
        if dev.type != "cuda":
            raise SevaB200Error("Seva parameters are on %s: seva_b200 runs on sm_100a only (no CPU fallback); "
                                "move the module with .to('cuda')" % dev)
        group = self._frame_group
        key = (dev.index, n, num_frames, h, w, ctx_len, None if group is None else id(group))
        eng = self._engines.get(key)
        if eng is None:
            ops.check_device()
            # workspaces / packed weights must be ordinary tensors even when the caller is inside
            # torch.inference_mode() (eval.py:1242): they are updated in place on every call
            with torch.inference_mode(False), torch.no_grad(), torch.cuda.device(dev):
                if self._packed is None:
                    self._packed = pack_weights(self)
                if len(self._engines) >= 2:  # plans own GBs of workspace: keep the two most recent shapes
                    self._engines.pop(next(iter(self._engines)))
                if group is None:
                    eng = Engine(self, self._packed, n, num_frames, h, w, ctx_len)
                else:
                    shard = FrameShard(group, num_frames)
                    eng = Engine(self, self._packed, (n // num_frames) * shard.T_local, shard.T_local, h, w, ctx_len,
                                 shard=shard)
            self._engines[key] = eng
        return eng

    @torch.compiler.disable  # a launch plan of hand-written kernels: opaque to torch.compile(MODEL) (demo.py:63)
    def forward(self, x: torch.Tensor, t: torch.Tensor, y: torch.Tensor, dense_y: torch.Tensor,
                num_frames: int | None = None, cond_epoch=None) -> torch.Tensor:
        """``cond_epoch`` (optional, from the guider's prepared conditioning): y / dense_y are unchanged since the
        last call with the same value, so the plan skips its conditioning-only launches."""
        num_frames = num_frames or self.params.num_frames
        if not x.is_cuda:
            raise SevaB200Error("Seva.forward needs CUDA tensors (seva_b200 has no CPU fallback)")
        assert y.ndim == 3
        n, cin, h, w = x.shape
        assert cin == self.params.in_channels and n % num_frames == 0
        dev = next(self.parameters()).device
        if x.device != dev:
            raise SevaB200Error(f"Seva.forward: input on {x.device}, parameters on {dev}")
        with torch.autocast("cuda", enabled=False), torch.cuda.device(dev):
            eng = self.engine(n, num_frames, h, w, y.shape[1])
            sh = eng.shard
            if sh is None:
                ops.pack_nchw(x.float(), eng.net_in)
                eng.bind(t, y, dense_y, epoch=cond_epoch)
                eng.run()
                out = ops.unpack_nchw(eng.eps, n, self.params.out_channels, h, w)
            else:  # this rank's frames in, every rank's frames out
                B = n // num_frames
                ops.pack_nchw(sh.local(x.float(), B), eng.net_in)
                if eng.cond_current(cond_epoch):
                    eng.bind(sh.local(t, B), None, None, epoch=cond_epoch)
                else:
                    eng.bind(sh.local(t, B), sh.local(y.float(), B), sh.local(dense_y.float(), B), ctx0=y[::num_frames],
                             epoch=cond_epoch)
                eng.run()
                loc = ops.unpack_nchw(eng.eps, eng.N, self.params.out_channels, h, w)
                out = torch.empty((n,) + tuple(loc.shape[1:]), device=loc.device, dtype=loc.dtype)
                sh.gather(out.view(n, -1), loc.view(eng.N, -1), B, 1)
        return out.to(x.dtype)


class SGMWrapper(nn.Module):
    def __init__(self, module: Seva):
        super().__init__()
        self.module = module

    @torch.compiler.disable  # nothing here for a tracing compiler: one cat, then the launch plan
    def forward(self, x: torch.Tensor, t: torch.Tensor, c: dict, **kwargs) -> torch.Tensor:
        concat = c.get("concat", None)
        if concat is not None and concat.numel() > 0:
            x = torch.cat((x, concat.type_as(x)), dim=1)
        if isinstance(self.module, Seva) and c.get("_cond_epoch") is not None:
            kwargs = dict(kwargs, cond_epoch=c["_cond_epoch"])
        return self.module(x, t=t, y=c["crossattn"], dense_y=c["dense_vector"], **kwargs)
