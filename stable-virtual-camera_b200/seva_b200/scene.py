"""Scene loop: the caller of the sampling path, as a multi-GPU schedule.

``run_one_scene`` (seva/eval.py:1324-1974) walks the planner's windows one after another on one GPU.  Here the same
windows form a DAG: first-pass / one-pass ``gt-*`` windows that condition on earlier outputs stay a chain, every other
window is independent (eval.py:1890-1906 reads only the anchor table), so

* independent windows are dealt round-robin to the ranks (``dist.schedule_chunks``), weights replicated, no data-path
  collective, one gather of the finished latents at the end (``dist.gather_chunks``);
* chained windows run on every rank in lockstep -- optionally model-parallel over ALL ranks (CFG halves x frame shards,
  dist.py) so the serial part of a two-pass run also scales -- and leave their result replicated.

The path works on latents.  The VAE and the CLIP conditioner (eval.py:1245-1248, 1316) are outside the hot path
(DESIGN.md section 6): they enter as the ``codec`` callables (identity / fixed projection in the tests and the bench).
Initial noise is drawn for ALL windows in window order from one generator on every rank, so an N-GPU run consumes the
random stream exactly like the sequential loop and returns the same latents as a 1-GPU run.
"""
from __future__ import annotations

from dataclasses import dataclass, field
from typing import Callable

import numpy as np
import torch

from . import planner as PL
from .geometry import get_plucker_coordinates, to_hom_pose


# ----------------------------------------------------------------------------------------------------
# per-window conditioning (eval.py:1152-1215 and 1245-1281)
# ----------------------------------------------------------------------------------------------------
def normalize_cameras(curr_c2ws: torch.Tensor, all_c2ws: torch.Tensor, camera_scale: float = 2.0):
    """Centre the window's cameras on the scene's (outlier-trimmed) mean position and scale translations so the
    first camera sits at distance ``camera_scale`` (eval.py:1173-1198).  Returns (c2w, w2c), both [T,4,4]."""
    c2w = to_hom_pose(curr_c2ws.float()).clone()
    pos = all_c2ws[:, :3, 3]
    d = torch.norm(pos - pos.median(0, keepdim=True).values, dim=-1)
    keep = d <= torch.clamp(torch.quantile(d, 0.97) * 10, max=1e6)
    c2w[:, :3, 3] -= pos[keep].mean(0, keepdim=True)
    w2c = torch.linalg.inv(c2w)
    first = torch.norm(c2w[0, :3, 3])
    scale = camera_scale if torch.isclose(first, torch.zeros(1, device=first.device), atol=1e-5).any() \
        else camera_scale / first
    w2c[:, :3, 3] *= scale
    c2w[:, :3, 3] *= scale
    return c2w, w2c


def window_cameras(curr_c2ws, curr_Ks, all_c2ws, hw, camera_scale: float = 2.0) -> dict:
    """c2w / K / Plucker map of one window, relative to its first frame (eval.py:1199-1213)."""
    c2w, w2c = normalize_cameras(curr_c2ws, all_c2ws, camera_scale)
    pl = get_plucker_coordinates(extrinsics_src=w2c[0], extrinsics=w2c, intrinsics=curr_Ks.float().clone(),
                                 target_size=hw)
    return dict(c2w=c2w, K=curr_Ks, plucker=pl)


def window_cond(latents: torch.Tensor, mask: torch.Tensor, token: torch.Tensor, plucker: torch.Tensor):
    """(cond, uncond) dicts of one window from the encoded input frames ``latents`` [n_in,4,h,w] (rows of the
    masked slots, in slot order), the pooled CLIP token [D] and the ray map [T,6,h,w] (eval.py:1245-1281)."""
    T, (h, w) = mask.shape[0], plucker.shape[-2:]
    lat1 = torch.nn.functional.pad(latents, (0, 0, 0, 0, 0, 1), value=1.0)
    replace = lat1.new_zeros(T, *lat1.shape[1:])
    replace[mask] = lat1
    cross = token[None, None].expand(T, 1, -1).contiguous()
    m = mask[:, None, None, None].to(plucker.dtype).expand(T, 1, h, w)
    c = {"crossattn": cross, "replace": replace, "concat": torch.cat([m, plucker], 1), "dense_vector": plucker}
    uc = {"crossattn": torch.zeros_like(cross), "replace": torch.zeros_like(replace),
          "concat": torch.cat([torch.zeros_like(m), plucker], 1), "dense_vector": plucker}
    return c, uc


# ----------------------------------------------------------------------------------------------------
# scene description and plan
# ----------------------------------------------------------------------------------------------------
@dataclass
class Scene:
    """Everything ``run_one_scene`` reads, in latent space.  ``input_*``: the M observed views; ``target_*``: the N
    views to generate (their order is the output order)."""
    input_latents: torch.Tensor     # [M,4,h,w]
    input_c2ws: torch.Tensor        # [M,4,4]
    input_Ks: torch.Tensor          # [M,3,3] normalised intrinsics
    target_c2ws: torch.Tensor       # [N,4,4]
    target_Ks: torch.Tensor         # [N,3,3]
    input_indices: list             # positions of the inputs in the scene's frame numbering
    target_indices: list            # positions of the targets
    prior_c2ws: torch.Tensor | None = None   # anchor cameras of a two-pass run [P,4,4]
    prior_Ks: torch.Tensor | None = None
    prior_indices: list | None = None

    @property
    def all_c2ws(self) -> torch.Tensor:
        n = len(self.input_indices) + len(self.target_indices)
        out = self.input_c2ws.new_zeros(n, 4, 4)
        out[self.input_indices] = to_hom_pose(self.input_c2ws.float())
        out[self.target_indices] = to_hom_pose(self.target_c2ws.float())
        return out


@dataclass
class Pass:
    T: int
    windows: list                   # planner.Window
    chained: bool                   # any window conditions on an earlier one
    cfg: float = 2.0


@dataclass
class ScenePlan:
    passes: list = field(default_factory=list)   # one entry (one-pass) or two (anchors, then targets)
    gt_input_inds: list = field(default_factory=list)
    prior_argsort: list = field(default_factory=list)


def plan_scene(scene: Scene, T, options: dict, task: str = "img2trajvid", use_traj_prior: bool = True) -> ScenePlan:
    """The reference's window lists for a scene (eval.py:1471-1495 one pass; 1655-1675 and 1835-1860 two passes)."""
    opt = dict(options, sampler_verbose=False)
    M = scene.input_c2ws.shape[0]
    cfg = options.get("cfg", 2.0)
    cfgs = list(cfg) if isinstance(cfg, (list, tuple)) else [cfg, cfg]
    plan = ScenePlan()
    if not use_traj_prior:
        T1 = T[0] if isinstance(T, (list, tuple)) else T
        r = PL.chunk_input_and_test(T1, scene.input_c2ws, scene.target_c2ws, scene.input_indices, scene.target_indices,
                                    options=opt, task=task, chunk_strategy=options.get("chunk_strategy", "gt"),
                                    gt_input_inds=list(range(M)))
        wins = PL.make_windows(T1, *r[1:], n_real_inputs=M, padding_mode=options.get("t_padding_mode", "last"))
        plan.passes.append(Pass(T1, wins, any(w.needs for w in wins), cfgs[0]))
        return plan
    assert scene.prior_c2ws is not None and scene.prior_indices is not None, \
        "`prior_c2ws` / `prior_indices` should be set when using 2-pass sampling."
    T1, T2 = (T[0], T[1]) if isinstance(T, (list, tuple)) else (T, T)
    r = PL.chunk_input_and_test(T1, scene.input_c2ws, scene.prior_c2ws, scene.input_indices, scene.prior_indices,
                                options=opt, task=task,
                                chunk_strategy=options.get("chunk_strategy_first_pass", "gt-nearest"),
                                gt_input_inds=list(range(M)))
    wins = PL.make_windows(T1, *r[1:], n_real_inputs=M, padding_mode=options.get("t_padding_mode", "last"))
    plan.passes.append(Pass(T1, wins, any(w.needs for w in wins), cfgs[0]))
    # anchors and inputs merged into one table sorted by scene position (eval.py:1812-1832)
    merged = list(scene.input_indices) + list(scene.prior_indices)
    order = np.argsort(merged).tolist()
    plan.prior_argsort = order
    plan.gt_input_inds = [order.index(i) for i in range(M)]
    src_c2ws = torch.cat([to_hom_pose(scene.input_c2ws.float()), to_hom_pose(scene.prior_c2ws.float())])[order]
    r = PL.chunk_input_and_test(T2, src_c2ws, scene.target_c2ws, np.array(merged)[order].tolist(),
                                scene.target_indices, options=opt, task=task,
                                chunk_strategy=options.get("chunk_strategy", "nearest"),
                                gt_input_inds=plan.gt_input_inds)
    wins = PL.make_windows(T2, *r[1:], n_real_inputs=None, padding_mode="last")
    plan.passes.append(Pass(T2, wins, False, cfgs[1] if len(cfgs) > 1 else cfgs[0]))
    return plan


# ----------------------------------------------------------------------------------------------------
# execution
# ----------------------------------------------------------------------------------------------------
@dataclass
class Codec:
    """Stand-ins for the modules either side of the path: ``encode`` images -> latents (VAE, eval.py:1246),
    ``decode`` latents -> images (eval.py:1316), ``embed`` images -> [n, D] tokens (CLIP, eval.py:1248).  The
    defaults keep everything in latent space: identity codec, and a fixed seeded projection of the pooled latent."""
    encode: Callable = lambda imgs: imgs
    decode: Callable = lambda lat: lat
    embed: Callable | None = None
    dim: int = 1024

    downsample: int = 1          # image pixels per latent pixel (8 for the VAE)
    latent_channels: int | None = None

    @classmethod
    def from_modules(cls, ae, conditioner) -> "Codec":
        """The modules ``do_sample`` calls either side of the path (eval.py:1245-1248, 1316): ``AutoEncoder.encode`` /
        ``.decode`` (seva_b200.autoencoder, N3) and the ``CLIPConditioner`` (seva_b200.conditioner, N4).  The scene's
        frame tables then hold RGB images in [-1, 1] and the sampler runs on their 8x smaller 4-channel latents."""
        return cls(encode=ae.encode, decode=ae.decode, embed=conditioner, dim=conditioner.params.embed_dim,
                   downsample=ae.downsample, latent_channels=4)

    def latent_shape(self, frame_shape: tuple) -> tuple:
        c, h, w = frame_shape
        return (self.latent_channels or c, h // self.downsample, w // self.downsample)

    def tokens(self, imgs: torch.Tensor) -> torch.Tensor:
        if self.embed is not None:
            return self.embed(imgs)
        g = torch.Generator().manual_seed(1234)
        proj = torch.randn(imgs.shape[1], self.dim, generator=g).to(imgs)
        return imgs.mean(dim=(-1, -2)) @ proj


@dataclass
class Rank:
    """Where this process sits: ``world`` ranks deal out independent windows; ``chain_guider`` / ``chain_frames``
    optionally make chained windows model-parallel over all ranks (dist.CFGSplit wrapper factory, frame group)."""
    rank: int = 0
    world: int = 1
    group: object = None
    chain_frames: object = None          # frame group for model.module.set_frame_group during chained windows
    chain_guider: Callable | None = None  # guider -> guider (e.g. lambda g: dist.CFGSplit(g, cfg_group, half))


def _window_inputs(win: PL.Window, src: dict, tgt: dict):
    """Assemble the window's T rows of every table (eval.py:1711-1747): latents/images, cameras, intrinsics."""
    return {k: PL.assemble(src[k][win.input_inds], tgt[k][win.test_inds], win.input_maps, win.test_maps)
            for k in ("img", "c2w", "K")}


def sample_window(model, denoiser, sampler, win: PL.Window, rows: dict, all_c2ws, x0, cfg: float, codec: Codec,
                  camera_scale: float = 2.0, device="cuda") -> torch.Tensor:
    """One sampling run (``do_sample``, eval.py:1218-1321): returns the decoded frames of the window's NEW slots,
    [len(test_sels), C, H, W], on the CPU like ``decode_output``."""
    T = len(win.input_maps)
    mask = torch.from_numpy(win.mask)
    imgs = rows["img"].to(device)
    hw = tuple(x0.shape[-2:])
    cam = window_cameras(rows["c2w"], rows["K"], all_c2ws, hw, camera_scale)
    with torch.inference_mode(), torch.autocast("cuda"):
        lat = codec.encode(imgs[mask.to(device)])
        token = codec.tokens(imgs[mask.to(device)]).mean(0)
        c, uc = window_cond(lat, mask.to(device), token, cam["plucker"].to(device))
        z = sampler(lambda inp, sigma, cc: denoiser(model, inp, sigma, cc, num_frames=T), x0.to(device), scale=cfg,
                    cond=c, uc=uc, verbose=False, c2w=cam["c2w"].to(device), K=cam["K"].to(device),
                    input_frame_mask=mask.to(device))
        out = codec.decode(z)
    return out.detach().cpu()[win.test_sels]


def window_noise_seed(base_seed: int, pass_index: int, window_index: int) -> int:
    """Seed of the per-step noise stream (``torch.randn_like`` in ``sampler_step``, sampling.py:361) of one window.
    The reference draws it from cuda:0's global Philox stream sequentially over the windows; here every window gets
    its own stream derived from (scene seed, pass, window), so a window's latents do not depend on which rank samples
    it or on what was sampled before it: an N-rank run returns the 1-rank run's latents."""
    return (int(base_seed) * 1000003 + pass_index * 10007 + window_index * 101 + 12345) % (2**63 - 1)


def _seed_step_noise(device, seed: int) -> None:
    dev = torch.device(device)
    if dev.type == "cuda":
        with torch.cuda.device(dev):
            torch.cuda.manual_seed(seed)
    else:
        torch.manual_seed(seed)


def run_scene(scene: Scene, plan: ScenePlan, model, denoiser, make_sampler: Callable, codec: Codec | None = None,
              where: Rank | None = None, generator: torch.Generator | None = None, camera_scale: float = 2.0,
              device="cuda", sample_fn: Callable | None = None) -> torch.Tensor:
    """Generate the scene's N target frames [N, C, H, W] (CPU tensor, target order).

    ``make_sampler(T, guider_wrap)`` returns the sampler for windows of T frames (``create_samplers``,
    eval.py:1092-1149); ``guider_wrap`` is None or the ``Rank.chain_guider`` to apply to its guider.
    ``sample_fn`` replaces ``sample_window`` (the CPU tests inject a stand-in; it receives the same arguments)."""
    from . import dist as D

    codec = codec or Codec()
    where = where or Rank()
    sample_fn = sample_fn or sample_window
    M = scene.input_c2ws.shape[0]
    all_c2ws = scene.all_c2ws
    frame_shape = tuple(scene.input_latents.shape[1:])   # latents, or RGB frames with a VAE codec
    lat_shape = codec.latent_shape(frame_shape)
    inputs = dict(img=scene.input_latents.cpu().float(), c2w=to_hom_pose(scene.input_c2ws.float()).cpu(),
                  K=scene.input_Ks.float().cpu())
    two_pass = len(plan.passes) == 2
    result = None
    base_seed = generator.initial_seed() if generator is not None else torch.initial_seed()
    for pi, ps in enumerate(plan.passes):
        last = pi == len(plan.passes) - 1
        if two_pass and not last:   # anchors
            table = dict(c2w=to_hom_pose(scene.prior_c2ws.float()).cpu(),
                         K=(scene.prior_Ks if scene.prior_Ks is not None
                            else scene.target_Ks[:1].repeat_interleave(scene.prior_c2ws.shape[0], 0)).float().cpu())
        else:
            table = dict(c2w=to_hom_pose(scene.target_c2ws.float()).cpu(), K=scene.target_Ks.float().cpu())
        table["img"] = torch.zeros((table["c2w"].shape[0],) + frame_shape)
        # every rank draws the noise of every window, in window order (torch.randn per do_sample call, eval.py:1290)
        noise = [torch.randn((ps.T,) + lat_shape, generator=generator) for _ in ps.windows]
        if two_pass and last:
            src = {k: torch.cat([inputs[k], anchors[k]])[plan.prior_argsort] for k in ("img", "c2w", "K")}
        done: dict = {}
        can_mp = where.world > 1 and (where.chain_frames is not None or where.chain_guider is not None)
        if where.world == 1 or ps.chained or (can_mp and 2 * len(ps.windows) <= where.world):
            # lockstep on every rank (model-parallel over all ranks when the caller set that up, else replicated);
            # later windows read [inputs; frames generated so far]
            par = can_mp
            if par and where.chain_frames is not None:
                model.module.set_frame_group(where.chain_frames)
            sampler = make_sampler(ps.T, where.chain_guider if par else None)
            gen_img, gen_idx = [], []
            for k, win in enumerate(ps.windows):
                if two_pass and last:
                    s = src
                else:
                    s = dict(img=torch.cat([inputs["img"]] + gen_img), c2w=torch.cat([inputs["c2w"], table["c2w"][gen_idx]]),
                             K=torch.cat([inputs["K"], table["K"][gen_idx]]))
                _seed_step_noise(device, window_noise_seed(base_seed, pi, k))
                done[k] = sample_fn(model, denoiser, sampler, win, _window_inputs(win, s, table), all_c2ws, noise[k],
                                    ps.cfg, codec, camera_scale, device)
                gen_img.append(done[k])
                gen_idx += win.test_inds
            if par and where.chain_frames is not None:
                model.module.set_frame_group(None)
        else:
            # independent windows: round-robin over the ranks, then one gather
            sampler = make_sampler(ps.T, None)
            mine = D.schedule_chunks(len(ps.windows), where.world)[where.rank]
            local = {}
            for k in mine:
                win = ps.windows[k]
                _seed_step_noise(device, window_noise_seed(base_seed, pi, k))
                out = sample_fn(model, denoiser, sampler, win, _window_inputs(win, src if (two_pass and last) else inputs, table),
                                all_c2ws, noise[k], ps.cfg, codec, camera_scale, device)
                pad = torch.zeros((ps.T,) + tuple(out.shape[1:]), dtype=out.dtype)  # equal shapes for the gather
                pad[: out.shape[0]] = out
                local[k] = pad.to(device) if torch.device(device).type == "cuda" else pad
            got = D.gather_chunks(local, len(ps.windows), where.group)
            done = {k: got[k][: len(ps.windows[k].test_inds)].cpu() for k in range(len(ps.windows))}
        order = [i for w in ps.windows for i in w.test_inds]
        frames = torch.cat([done[k] for k in range(len(ps.windows))])
        if two_pass and not last:
            anchors = dict(img=frames, c2w=table["c2w"], K=table["K"])  # generation order == anchor order
        else:
            result = frames[np.argsort(order)]   # eval.py:1957-1959
    return result
