"""ORACLE — test infrastructure only (never imported by the product path).

Plain PyTorch fp32 restatement of the reference algorithm for the Seva denoising step, written functionally over
a state dict so it needs neither the reference package nor our kernels.  Every function cites the reference
file:line it follows (paths relative to the reference repo root).  Pinning: ``oracle/gen_golden.py`` runs the
real reference (imported from /root/reference in the build container) on seeded inputs and commits its outputs
under ``tests/golden/``; ``tests/test_oracle_cpu.py`` checks this restatement against those vectors, and the GPU
parity tests then compare the CUDA path with this restatement (and with the golden vectors directly).
The reference ships no tests or golden vectors of its own, so these self-generated vectors
are the only pin available.
"""
from __future__ import annotations

import math

import numpy as np
import torch
import torch.nn.functional as F

# ---------------------------------------------------------------------------------------------------
# network: seva/model.py:176-216, seva/modules/layers.py, seva/modules/transformer.py
# ---------------------------------------------------------------------------------------------------


def timestep_embedding(t: torch.Tensor, dim: int, max_period: int = 10000) -> torch.Tensor:
    """layers.py:11-32"""
    half = dim // 2
    freqs = torch.exp(-math.log(max_period) * torch.arange(0, half, dtype=torch.float32) / half).to(t.device)
    args = t[:, None].float() * freqs[None]
    emb = torch.cat([torch.cos(args), torch.sin(args)], dim=-1)
    if dim % 2:
        emb = torch.cat([emb, torch.zeros_like(emb[:, :1])], dim=-1)
    return emb


def _lin(sd, k, x):
    return F.linear(x, sd[k + ".weight"], sd.get(k + ".bias"))


def _attention(sd, k, x, context, heads, kv_gather=None):
    """transformer.py:59-74.  ``kv_gather`` (frame sequence-parallel check only): maps this rank's K / V
    [batch, local keys, C] to the keys of all frames."""
    q = _lin(sd, k + ".to_q", x)
    ctx = x if context is None else context
    kk, v = _lin(sd, k + ".to_k", ctx), _lin(sd, k + ".to_v", ctx)
    if kv_gather is not None:
        kk, v = kv_gather(kk), kv_gather(v)
    sp = lambda t: t.reshape(t.shape[0], t.shape[1], heads, -1).transpose(1, 2)
    o = F.scaled_dot_product_attention(sp(q), sp(kk), sp(v))
    o = o.transpose(1, 2).reshape(x.shape[0], x.shape[1], -1)
    return _lin(sd, k + ".to_out.0", o)


def _ff(sd, k, x):
    """transformer.py:8-34: GEGLU (exact GELU) then Linear"""
    a, gate = _lin(sd, k + ".net.0.proj", x).chunk(2, dim=-1)
    return _lin(sd, k + ".net.2", a * F.gelu(gate))


def _ln(sd, k, x):
    return F.layer_norm(x, (x.shape[-1],), sd[k + ".weight"], sd[k + ".bias"], 1e-5)


def resblock(sd, k, x, emb, dense):
    """layers.py:120-139"""
    h = F.silu(F.group_norm(x, 32, sd[k + ".in_layers.0.weight"], sd[k + ".in_layers.0.bias"], 1e-5))
    d = F.interpolate(dense, size=h.shape[2:], mode="bilinear", align_corners=True)
    d = F.conv2d(d, sd[k + ".dense_emb_layers.0.weight"], sd[k + ".dense_emb_layers.0.bias"])
    scale, shift = torch.chunk(d, 2, dim=1)
    h = h * (1 + scale) + shift
    h = F.conv2d(h, sd[k + ".in_layers.2.weight"], sd[k + ".in_layers.2.bias"], padding=1)
    e = F.linear(F.silu(emb), sd[k + ".emb_layers.1.weight"], sd[k + ".emb_layers.1.bias"])
    h = h + e[:, :, None, None]
    h = F.silu(F.group_norm(h, 32, sd[k + ".out_layers.0.weight"], sd[k + ".out_layers.0.bias"], 1e-5))
    h = F.conv2d(h, sd[k + ".out_layers.3.weight"], sd[k + ".out_layers.3.bias"], padding=1)
    if k + ".skip_connection.weight" in sd:
        x = F.conv2d(x, sd[k + ".skip_connection.weight"], sd[k + ".skip_connection.bias"])
    return x + h


def multiview_transformer(sd, k, x, context, T, heads, unflatten, sp=None):
    """transformer.py:215-247 (any transformer_depth: the state dict decides).  With ``sp`` (frame sequence-parallel restatement) ``x`` and
    ``context`` hold this rank's T frames of every batch element, ``sp.ctx0`` is ``context[::T_total]`` of the
    full batch and ``sp.gather(t, per_frame)`` concatenates dim 1 of ``t`` over the ranks (per_frame = keys per
    frame): only the 3D and the time-mix self-attention see other ranks' frames."""
    n, c, h, w = x.shape
    x_in = x
    ctx0 = context[::T] if sp is None else sp.ctx0
    g3 = None if sp is None else (lambda t: sp.gather(t, h * w))
    gt = None if sp is None else (lambda t: sp.gather(t, 1))
    time_ctx = ctx0.repeat_interleave(h * w, dim=0)  # "b ... -> (b n) ..." (:222-226)
    if unflatten:
        context = ctx0
    x = F.group_norm(x, 32, sd[k + ".norm.weight"], sd[k + ".norm.bias"], 1e-6)
    x = x.permute(0, 2, 3, 1).reshape(n, h * w, c)
    x = _lin(sd, k + ".proj_in", x)
    d = 0
    while f"{k}.transformer_blocks.{d}.norm1.weight" in sd:   # transformer_depth pairs (:234-241)
        tb, tm = f"{k}.transformer_blocks.{d}", f"{k}.time_mix_blocks.{d}"
        if unflatten:
            x = x.reshape(n // T, T * h * w, c)
        # TransformerBlock.forward (:106-110)
        x = _attention(sd, tb + ".attn1", _ln(sd, tb + ".norm1", x), None, heads, g3 if unflatten else None) + x
        x = _attention(sd, tb + ".attn2", _ln(sd, tb + ".norm2", x), context, heads) + x
        x = _ff(sd, tb + ".ff", _ln(sd, tb + ".norm3", x)) + x
        if unflatten:
            x = x.reshape(n, h * w, c)
        # TransformerBlockTimeMix.forward (:145-155)
        s = h * w
        y = x.reshape(n // T, T, s, c).transpose(1, 2).reshape(-1, T, c)
        y = _ff(sd, tm + ".ff_in", _ln(sd, tm + ".norm_in", y)) + y
        y = _attention(sd, tm + ".attn1", _ln(sd, tm + ".norm1", y), None, heads, gt) + y
        y = _attention(sd, tm + ".attn2", _ln(sd, tm + ".norm2", y), time_ctx, heads) + y
        y = _ff(sd, tm + ".ff", _ln(sd, tm + ".norm3", y))
        y = y.reshape(n // T, s, T, c).transpose(1, 2).reshape(n, s, c)
        x = x + y  # SkipConnect (:158-165)
        d += 1
    x = _lin(sd, k + ".proj_out", x)
    return x.reshape(n, h, w, c).permute(0, 3, 1, 2) + x_in


def block_layout(params):
    """The block list seva/model.py:54-174 builds, as (kind, ...) tuples per block."""
    mc = params.model_channels
    inp = [[("conv_in",)]]
    ch, ds = mc, 1
    last = len(params.channel_mult) - 1
    for level, mult in enumerate(params.channel_mult):
        for _ in range(params.num_res_blocks):
            blk = [("res",)]
            ch = mult * mc
            if ds in params.attention_resolutions:
                blk.append(("mvt", ch // params.num_head_channels, f"input_ds{ds}"))
            inp.append(blk)
        if level != last:
            ds *= 2
            inp.append([("down",)])
    mid = [("res",), ("mvt", ch // params.num_head_channels, f"middle_ds{ds}"), ("res",)]
    out = []
    for level in range(last, -1, -1):
        mult = params.channel_mult[level]
        for i in range(params.num_res_blocks + 1):
            blk = [("res",)]
            ch = mc * mult
            if ds in params.attention_resolutions:
                blk.append(("mvt", ch // params.num_head_channels, f"output_ds{ds}"))
            if level and i == params.num_res_blocks:
                ds //= 2
                blk.append(("up",))
            out.append(blk)
    return inp, mid, out


def seva_forward(sd, params, x, t, y, dense_y, num_frames=None, sp=None):
    """Seva.forward, model.py:176-216 (fp32).  ``sp``: see multiview_transformer (num_frames = local frames)."""
    T = num_frames or params.num_frames
    emb = timestep_embedding(t, params.model_channels)
    emb = _lin(sd, "time_embed.2", F.silu(_lin(sd, "time_embed.0", emb)))
    inp, mid, out = block_layout(params)

    def run(prefix, blk, h):
        for j, layer in enumerate(blk):
            k = f"{prefix}.{j}"
            if layer[0] == "conv_in":
                h = F.conv2d(h, sd[k + ".weight"], sd[k + ".bias"], padding=1)
            elif layer[0] == "res":
                h = resblock(sd, k, h, emb, dense_y)
            elif layer[0] == "mvt":
                h = multiview_transformer(sd, k, h, y, T, layer[1], layer[2] in params.unflatten_names, sp)
            elif layer[0] == "down":
                h = F.conv2d(h, sd[k + ".op.weight"], sd[k + ".op.bias"], stride=2, padding=1)
            elif layer[0] == "up":
                h = F.interpolate(h, scale_factor=2, mode="nearest")
                h = F.conv2d(h, sd[k + ".conv.weight"], sd[k + ".conv.bias"], padding=1)
        return h

    hs, h = [], x
    for i, blk in enumerate(inp):
        h = run(f"input_blocks.{i}", blk, h)
        hs.append(h)
    h = run("middle_block", mid, h)
    for i, blk in enumerate(out):
        h = run(f"output_blocks.{i}", blk, torch.cat([h, hs.pop()], dim=1))
    h = F.silu(F.group_norm(h, 32, sd["out.0.weight"], sd["out.0.bias"], 1e-5))
    return F.conv2d(h, sd["out.2.weight"], sd["out.2.bias"], padding=1)


def sgm_forward(sd, params, x, t, c, num_frames=None):
    """SGMWrapper.forward, model.py:224-234"""
    x = torch.cat((x, c["concat"]), dim=1)
    return seva_forward(sd, params, x, t, c["crossattn"], c["dense_vector"], num_frames)


# ---------------------------------------------------------------------------------------------------
# sampling: seva/sampling.py
# ---------------------------------------------------------------------------------------------------
def ddpm_sigmas(n, num_timesteps=1000, linear_start=5e-6, linear_end=0.012, log_snr_shift=2.4):
    """sampling.py:28-102: scaled-linear betas (fp64) -> sigma table, subsampled, flipped (descending)."""
    betas = (torch.linspace(linear_start**0.5, linear_end**0.5, num_timesteps, dtype=torch.float64) ** 2).numpy()
    acp = np.cumprod(1.0 - betas, axis=0)
    if n < num_timesteps:
        steps = np.linspace(num_timesteps - 1, 0, n, endpoint=False).astype(int)[::-1]
        acp = acp[steps]
    elif n != num_timesteps:
        raise ValueError(n)
    sig = ((1 - acp) / acp) ** 0.5 * np.exp(log_snr_shift)
    return torch.flip(torch.tensor(sig, dtype=torch.float32), (0,))


def denoiser_table(num_idx=1000):
    """DiscreteDenoiser.register_sigmas, sampling.py:121-124 (ascending)."""
    return torch.flip(ddpm_sigmas(num_idx), (0,))


def sigma_to_idx(table, sigma):
    """sampling.py:126-128"""
    return (sigma - table[:, None]).abs().argmin(dim=0).view(sigma.shape)


def denoise(net, table, inp, sigma, cond):
    """DiscreteDenoiser.__call__, sampling.py:133-152; ``net(x, idx, cond)`` is the network."""
    cond = dict(cond)
    sigma = table[sigma_to_idx(table, sigma)]
    shape = sigma.shape
    s = sigma.reshape(-1, *([1] * (inp.ndim - 1)))
    c_skip, c_out, c_in = torch.ones_like(s), -s, 1 / (s**2 + 1.0) ** 0.5
    c_noise = sigma_to_idx(table, s.reshape(shape).clone())
    if "replace" in cond:
        r, mask = cond.pop("replace").split((inp.shape[1], 1), dim=1)
        inp = inp * (1 - mask) + r * mask
    return net(inp * c_in, c_noise, cond) * c_out + inp * c_skip


def camera_dist(a, b, mode):
    """geometry.py:12-40"""
    if mode == "rotation":
        m = torch.matmul(a[:, None, :3, :3], b[None, :, :3, :3].transpose(-1, -2))
        tr = m.diagonal(offset=0, dim1=-2, dim2=-1).sum(-1)
        return torch.acos(((tr - 1) / 2).clamp(-1, 1)) * (180 / torch.pi)
    return torch.norm(a[:, None, :3, 3] - b[None, :, :3, 3], dim=-1)


def close_frame(c2w, K, mask):
    """MultiviewScaleRule, sampling.py:171-180"""
    ci = c2w[mask]
    rot = camera_dist(c2w, ci, "rotation").min(-1).values
    tr = camera_dist(c2w, ci, "translation").min(-1).values
    kd = ((K[:, None] - K[mask][None]).flatten(-2) == 0).all(-1).any(-1)
    return (rot < 10.0) & (tr < 1e-5) & kd


def multiview_scale(scale, cfg_min, c2w, K, mask):
    """MultiviewScaleRule.__call__, sampling.py:164-187 -> per-frame scale"""
    close = close_frame(c2w, K, mask)
    if isinstance(scale, torch.Tensor):
        scale = scale.clone()
        scale[close] = cfg_min
        return scale
    return torch.where(close, cfg_min, scale)


def temporal_scale(scale, cfg_min, mask, T):
    """MultiviewTemporalCFG.__call__, sampling.py:286-296 (before the MultiviewCFG rule)."""
    ar = torch.arange(T)
    dm = (ar[None] - ar[:, None]).abs()
    m = mask.reshape(-1, T)
    md = (dm[None].to(m.device) + (~m[:, None]) * T).min(-1)[0]
    md = md / md.max(-1, keepdim=True)[0].clamp(min=1)
    return (md * (scale - cfg_min) + cfg_min).reshape(-1)


def prepare_inputs(x, s, c, uc):
    """VanillaCFG.prepare_inputs, sampling.py:231-242 (uncond first)"""
    return torch.cat([x] * 2), torch.cat([s] * 2), {k: torch.cat((uc[k], c[k]), 0) for k in c}


def sampler_step(net, table, x, sigma, next_sigma, frame_scale, c, uc, eps, gamma=0.0, s_noise=1.0):
    """EulerEDMSampler.sampler_step, sampling.py:347-368 with the noise draw ``eps`` injected."""
    sigma_hat = sigma * (gamma + 1.0) + 1e-6
    bc = lambda v: v.reshape(-1, *([1] * (x.ndim - 1)))
    x = x + eps * s_noise * bc(sigma_hat**2 - sigma**2) ** 0.5
    den = denoise(net, table, *prepare_inputs(x, sigma_hat, c, uc))
    u, cc = den.chunk(2)
    den = u + bc(frame_scale) * (cc - u)  # ConstantGuidance, sampling.py:213
    d = (x - den) / bc(sigma_hat)
    return x + bc(next_sigma - sigma_hat) * d, den


def sample(net, table, x, frame_scale, c, uc, num_steps, eps_all):
    """EulerEDMSampler.__call__, sampling.py:370-405 with s_churn = 0 (eval.py:1141); x is not mutated."""
    sig = torch.cat([ddpm_sigmas(num_steps), torch.zeros(1)]).to(x.device)
    x = x * torch.sqrt(1.0 + sig[0] ** 2.0)
    ones = x.new_ones([x.shape[0]])
    for i in range(num_steps):
        x, _ = sampler_step(net, table, x, ones * sig[i], ones * sig[i + 1], frame_scale, c, uc, eps_all[i])
    return x


# ---------------------------------------------------------------------------------------------------
# geometry: seva/geometry.py:43-165
# ---------------------------------------------------------------------------------------------------
def default_intrinsics(fov_rad=0.9424777960769379):
    """geometry.py:58-79, aspect ratio 1"""
    f = 0.5 / math.tan(0.5 * fov_rad)
    return torch.tensor([[[f, 0.0, 0.5], [0.0, f, 0.5], [0.0, 0.0, 1.0]]], dtype=torch.float32)


def plucker(w2c_src, w2c, K_norm, h, w):
    """get_plucker_coordinates, geometry.py:119-165 (K_norm normalised; not mutated here)."""
    K = K_norm.clone().float()
    K[:, :2] *= torch.tensor([w, h], dtype=torch.float32).view(1, -1, 1)
    rel = w2c @ torch.linalg.inv(w2c_src)[None]
    ys, xs = torch.meshgrid(torch.arange(h, dtype=torch.float32) + 0.5, torch.arange(w, dtype=torch.float32) + 0.5,
                            indexing="ij")
    grid = torch.stack([xs, ys, torch.ones_like(xs)], dim=-1).view(-1, 3)          # [HW,3]
    cam = grid @ torch.linalg.inv(K).transpose(-1, -2)                            # [T,HW,3]
    pose = rel[:, :3, :]
    hom = torch.eye(4)[None].repeat(pose.shape[0], 1, 1)
    hom[:, :3, :] = pose
    pinv = torch.linalg.inv(hom)[:, :3, :4]
    to_world = lambda X: torch.cat([X, torch.ones_like(X[..., :1])], -1) @ pinv.transpose(-1, -2)
    pts, ctr = to_world(cam.expand(pose.shape[0], -1, -1)), to_world(torch.zeros_like(cam).expand(pose.shape[0], -1, -1))
    rays = F.normalize(pts - ctr, dim=-1)
    pl = torch.cat((rays, torch.cross(ctr, rays, dim=-1)), dim=-1)
    return pl.permute(0, 2, 1).reshape(pose.shape[0], 6, h, w)
