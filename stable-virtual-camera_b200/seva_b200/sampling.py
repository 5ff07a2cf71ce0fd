"""Sampler / denoiser / guidance behind the reference's ``seva.sampling`` surface (seva/sampling.py:1-405).

Same classes, argument meaning and error behaviour: ``DDPMDiscretization`` (:57-102), ``DiscreteDenoiser``
(:105-152), ``VanillaCFG`` / ``MultiviewCFG`` / ``MultiviewTemporalCFG`` (:216-298), ``EulerEDMSampler``
(:301-405) with public ``prepare_sampling_loop`` / ``get_sigma_gen`` / ``sampler_step`` (the Gradio sampler
subclass calls them, eval.py:1060-1082).  On CUDA tensors the per-step arithmetic runs in our kernels:
``seva_sigma_quantize``, ``seva_add_noise``, ``seva_denoise_prepare`` -> network plan -> ``seva_denoise_finish``,
``seva_multiview_scale`` and the fused ``seva_cfg_euler_step``.  The sigma tables are host fp64 -> fp32 exactly
as the reference builds them, so sigma indices are bit-exact.
"""
from __future__ import annotations

import itertools

import numpy as np
import torch
import torch.nn as nn

from . import ops
from .geometry import get_camera_dist

try:  # progress bars are cosmetic
    from tqdm import tqdm
except Exception:  # pragma: no cover
    tqdm = None


def append_dims(x: torch.Tensor, target_dims: int) -> torch.Tensor:
    dims_to_append = target_dims - x.ndim
    if dims_to_append < 0:
        raise ValueError(f"input has {x.ndim} dims but target_dims is {target_dims}, which is less")
    return x[(...,) + (None,) * dims_to_append]


def append_zero(x: torch.Tensor) -> torch.Tensor:
    return torch.cat([x, x.new_zeros([1])])


def to_d(x: torch.Tensor, sigma: torch.Tensor, denoised: torch.Tensor) -> torch.Tensor:
    return (x - denoised) / append_dims(sigma, x.ndim)


def make_betas(num_timesteps: int, linear_start: float = 1e-4, linear_end: float = 2e-2) -> np.ndarray:
    betas = torch.linspace(linear_start**0.5, linear_end**0.5, num_timesteps, dtype=torch.float64) ** 2
    return betas.numpy()


def generate_roughly_equally_spaced_steps(num_substeps: int, max_step: int) -> np.ndarray:
    return np.linspace(max_step - 1, 0, num_substeps, endpoint=False).astype(int)[::-1]


class EpsScaling(object):
    def __call__(self, sigma: torch.Tensor):
        c_skip = torch.ones_like(sigma, device=sigma.device)
        c_out = -sigma
        c_in = 1 / (sigma**2 + 1.0) ** 0.5
        c_noise = sigma.clone()
        return c_skip, c_out, c_in, c_noise


class DDPMDiscretization(object):
    def __init__(self, linear_start: float = 5e-06, linear_end: float = 0.012, num_timesteps: int = 1000,
                 log_snr_shift: float | None = 2.4):
        self.num_timesteps = num_timesteps
        betas = make_betas(num_timesteps, linear_start=linear_start, linear_end=linear_end)
        self.log_snr_shift = log_snr_shift
        self.alphas_cumprod = np.cumprod(1.0 - betas, axis=0)

    def get_sigmas(self, n: int, device: str | torch.device = "cpu") -> torch.Tensor:
        if n < self.num_timesteps:
            acp = self.alphas_cumprod[generate_roughly_equally_spaced_steps(n, self.num_timesteps)]
        elif n == self.num_timesteps:
            acp = self.alphas_cumprod
        else:
            raise ValueError(f"Expected n <= {self.num_timesteps}, but got n = {n}.")
        sigmas = ((1 - acp) / acp) ** 0.5
        if self.log_snr_shift is not None:
            sigmas = sigmas * np.exp(self.log_snr_shift)
        return torch.flip(torch.tensor(sigmas, dtype=torch.float32, device=device), (0,))

    def __call__(self, n: int, do_append_zero: bool = True, flip: bool = False,
                 device: str | torch.device = "cpu") -> torch.Tensor:
        sigmas = self.get_sigmas(n, device=device)
        sigmas = append_zero(sigmas) if do_append_zero else sigmas
        return sigmas if not flip else torch.flip(sigmas, (0,))


def _is_fused_network(network) -> bool:
    from .model import SGMWrapper, Seva

    network = getattr(network, "_orig_mod", network)  # torch.compile(MODEL) wrapper (demo.py:63)
    return isinstance(network, SGMWrapper) and isinstance(network.module, Seva)


class DiscreteDenoiser(object):
    sigmas: torch.Tensor

    def __init__(self, discretization: DDPMDiscretization, num_idx: int = 1000, device: str | torch.device = "cpu"):
        self.scaling = EpsScaling()
        self.discretization = discretization
        self.num_idx = num_idx
        self.device = device
        self.register_sigmas()

    def register_sigmas(self):
        self.sigmas = self.discretization(self.num_idx, do_append_zero=False, flip=True, device=self.device)

    def sigma_to_idx(self, sigma: torch.Tensor) -> torch.Tensor:
        if sigma.is_cuda and self.sigmas.is_cuda and sigma.dtype == torch.float32:
            return ops.sigma_quantize(sigma.reshape(-1), self.sigmas)[0].view(sigma.shape)
        dists = sigma - self.sigmas[:, None]
        return dists.abs().argmin(dim=0).view(sigma.shape)

    def idx_to_sigma(self, idx: torch.Tensor | int) -> torch.Tensor:
        return self.sigmas[idx]

    def __call__(self, network: nn.Module, input: torch.Tensor, sigma: torch.Tensor, cond: dict,
                 **additional_model_inputs) -> torch.Tensor:
        if self._fusable(network, input, sigma, cond):
            return self._call_fused(network, input, sigma, cond, **additional_model_inputs)
        sigma = self.idx_to_sigma(self.sigma_to_idx(sigma))
        sigma_shape = sigma.shape
        sigma = append_dims(sigma, input.ndim)
        c_skip, c_out, c_in, c_noise = self.scaling(sigma)
        c_noise = self.sigma_to_idx(c_noise.reshape(sigma_shape))
        if "replace" in cond:
            x, mask = cond.pop("replace").split((input.shape[1], 1), dim=1)
            input = input * (1 - mask) + x * mask
        return network(input * c_in, c_noise, cond, **additional_model_inputs) * c_out + input * c_skip

    def _fusable(self, network, input, sigma, cond) -> bool:
        """The fused kernels hard-code the released model's boundary: 4 latent + 7 concat (mask, Plucker) input
        channels, a 5-channel ``replace`` (latent + mask), 4 output channels, everything resident on the model's
        GPU.  Anything else takes the generic branch below, which states the reference's arithmetic in PyTorch and
        raises the reference's shape errors."""
        if not (input.is_cuda and self.sigmas.is_cuda and sigma.is_cuda and _is_fused_network(network)):
            return False
        module = getattr(network, "_orig_mod", network).module
        p = module.params
        if next(module.parameters()).device != input.device or self.sigmas.device != input.device:
            return False
        if not (input.ndim == 4 and input.shape[1] == 4 and p.in_channels == 11 and p.out_channels == 4):
            return False
        cc, rep = cond.get("concat"), cond.get("replace")
        same = lambda t: isinstance(t, torch.Tensor) and t.device == input.device
        if not (same(cc) and cc.ndim == 4 and cc.shape[1] == 7 and cc.shape[0] == input.shape[0]
                and cc.shape[2:] == input.shape[2:]):
            return False
        if rep is not None and not (same(rep) and rep.ndim == 4 and rep.shape[1] == 5 and rep.shape[0] == input.shape[0]
                                    and rep.shape[2:] == input.shape[2:]):
            return False
        return same(cond.get("crossattn")) and same(cond.get("dense_vector")) and cond["crossattn"].ndim == 3 \
            and cond["dense_vector"].shape[1] == p.dense_in_channels and sigma.numel() == input.shape[0]

    def _call_fused(self, network, input, sigma, cond, num_frames=None, **extra):
        """Same arithmetic as above with the elementwise work fused around the network plan."""
        module = getattr(network, "_orig_mod", network).module
        num_frames = num_frames or module.params.num_frames
        n, _, h, w = input.shape
        with torch.autocast("cuda", enabled=False), torch.cuda.device(input.device):
            idx, _, c_in, c_out = ops.sigma_quantize(sigma.float().reshape(-1), self.sigmas)
            replace = cond.pop("replace", None)  # the reference pops it from the per-step dict (sampling.py:147)
            eng = module.engine(n, num_frames, h, w, cond["crossattn"].shape[1])
            sh = eng.shard
            if sh is not None:
                return self._call_fused_sharded(eng, sh, input, replace, cond, idx, c_in, c_out, num_frames)
            blended = torch.empty((n, 4, h, w), device=input.device, dtype=torch.float32)
            ops.denoise_prepare(input.float().contiguous(), None if replace is None else replace.float().contiguous(),
                                cond["concat"].float().contiguous(), c_in, eng.net_in, blended, n, n, h, w)
            eng.bind(idx, cond["crossattn"], cond["dense_vector"], epoch=cond.get(COND_EPOCH), t_range=self.sigmas.numel())
            eng.run()
            out = torch.empty_like(blended)
            ops.denoise_finish(eng.eps, blended, c_out, out, n, h, w)
        return out

    def _call_fused_sharded(self, eng, sh, input, replace, cond, idx, c_in, c_out, num_frames):
        """Frame sequence-parallel form of ``_call_fused``: the same kernels on this rank's frames, then one gather of
        the denoised frames so that every rank returns the full batch (x stays replicated in the sampler)."""
        n, _, h, w = input.shape
        B, nl = n // num_frames, eng.N
        loc = lambda t: sh.local(t.float(), B)
        blended = torch.empty((nl, 4, h, w), device=input.device, dtype=torch.float32)
        ops.denoise_prepare(loc(input), None if replace is None else loc(replace), loc(cond["concat"]), loc(c_in),
                            eng.net_in, blended, nl, nl, h, w)
        epoch = cond.get(COND_EPOCH)
        if eng.cond_current(epoch):
            eng.bind(sh.local(idx, B), None, None, epoch=epoch, t_range=self.sigmas.numel())
        else:
            ctx = cond["crossattn"].float()
            eng.bind(sh.local(idx, B), sh.local(ctx, B), loc(cond["dense_vector"]), ctx0=ctx[::num_frames], epoch=epoch,
                     t_range=self.sigmas.numel())
        eng.run()
        out_l = torch.empty_like(blended)
        ops.denoise_finish(eng.eps, blended, loc(c_out), out_l, nl, h, w)
        out = torch.empty((n, 4, h, w), device=input.device, dtype=torch.float32)
        sh.gather(out.view(n, -1), out_l.view(nl, -1), B, 1)
        return out


class ConstantScaleRule(object):
    def __call__(self, scale: float | torch.Tensor) -> float | torch.Tensor:
        return scale


class MultiviewScaleRule(object):
    def __init__(self, min_scale: float = 1.0):
        self.min_scale = min_scale

    def close_frame(self, c2w: torch.Tensor, K: torch.Tensor, input_frame_mask: torch.Tensor) -> torch.Tensor:
        """bool [T]: frame shares pose (rot < 10 deg, |dt| < 1e-5) and intrinsics with some input frame."""
        if c2w.is_cuda and c2w.shape[-2:] == (4, 4):
            return ops.multiview_scale(c2w.float(), K.float(), input_frame_mask, 0.0, self.min_scale)[1]
        c2w_input = c2w[input_frame_mask]
        rotation_diff = get_camera_dist(c2w, c2w_input, mode="rotation").min(-1).values
        translation_diff = get_camera_dist(c2w, c2w_input, mode="translation").min(-1).values
        K_diff = ((K[:, None] - K[input_frame_mask][None]).flatten(-2) == 0).all(-1).any(-1)
        return (rotation_diff < 10.0) & (translation_diff < 1e-5) & K_diff

    def __call__(self, scale: float | torch.Tensor, c2w: torch.Tensor, K: torch.Tensor,
                 input_frame_mask: torch.Tensor) -> torch.Tensor:
        close = self.close_frame(c2w, K, input_frame_mask)
        if isinstance(scale, torch.Tensor):
            scale = scale.clone()
            scale[close] = self.min_scale
        elif isinstance(scale, float):
            scale = torch.where(close, self.min_scale, scale)
        else:
            raise ValueError(f"Invalid scale type {type(scale)}.")
        return scale


class ConstantScaleSchedule(object):
    def __call__(self, sigma: float | torch.Tensor, scale: float | torch.Tensor) -> float | torch.Tensor:
        if isinstance(sigma, float):
            return scale
        elif isinstance(sigma, torch.Tensor):
            if len(sigma.shape) == 1 and isinstance(scale, torch.Tensor):
                sigma = append_dims(sigma, scale.ndim)
            return scale * torch.ones_like(sigma)
        else:
            raise ValueError(f"Invalid sigma type {type(sigma)}.")


class ConstantGuidance(object):
    def __call__(self, uncond: torch.Tensor, cond: torch.Tensor, scale: float | torch.Tensor) -> torch.Tensor:
        if isinstance(scale, torch.Tensor) and len(scale.shape) == 1:
            scale = append_dims(scale, cond.ndim)
        return uncond + scale * (cond - uncond)


_CAT_KEYS = ("vector", "crossattn", "concat", "replace", "dense_vector")  # seva/sampling.py:234
COND_EPOCH = "_cond_epoch"  # private key of the prepared conditioning dict (ignored by the reference's SGMWrapper)
_EPOCHS = itertools.count(1)


def _version(t) -> int:
    try:
        return t._version
    except RuntimeError:  # inference tensors do not track versions
        return -1


def _same_cond(a: list, b: list) -> bool:
    return len(a) == len(b) and all(p[0] == q[0] and p[1] is q[1] and p[2] is q[2] and p[3:] == q[3:]
                                    for p, q in zip(a, b))


class VanillaCFG(object):
    def __init__(self):
        self.scale_rule = ConstantScaleRule()
        self.scale_schedule = ConstantScaleSchedule()
        self.guidance = ConstantGuidance()

    def __call__(self, x: torch.Tensor, sigma: float | torch.Tensor, scale: float | torch.Tensor) -> torch.Tensor:
        x_u, x_c = x.chunk(2)
        scale = self.scale_rule(scale)
        scale_value = self.scale_schedule(sigma, scale)
        return self.guidance(x_u, x_c, scale_value)

    def prepare_inputs(self, x: torch.Tensor, s: torch.Tensor, c: dict, uc: dict):
        """[uncond; cond] stacking (sampling.py:231-242).  The stacked conditioning does not change inside a
        sampling loop, so it is built once per (c, uc) and loop -- the sampler drops the cache in
        ``prepare_sampling_loop`` -- and tagged with a ``_cond_epoch`` the denoiser hands to the launch plan,
        which then skips its conditioning-only launches.  A fresh dict is returned every call: the
        denoiser pops ``replace`` from it (sampling.py:147)."""
        ident = [(k, c[k], uc[k], _version(c[k]), _version(uc[k])) for k in c if k in _CAT_KEYS]
        cache = self.__dict__.get("_cond_cache")
        if cache is None or not _same_cond(cache[0], ident):
            c_out = {k: torch.cat((u, v), 0) for k, v, u, _, _ in ident}
            c_out[COND_EPOCH] = next(_EPOCHS)
            cache = self.__dict__["_cond_cache"] = (ident, c_out)  # holds c / uc alive: identities stay unique
        c_out = dict(cache[1])
        for k in c:
            if k not in _CAT_KEYS:
                assert c[k] == uc[k]
                c_out[k] = c[k]
        return torch.cat([x] * 2), torch.cat([s] * 2), c_out

    def reset_cache(self) -> None:
        self.__dict__.pop("_cond_cache", None)

    def frame_scale(self, x: torch.Tensor, sigma: torch.Tensor, scale, **guider_kwargs) -> torch.Tensor | None:
        """Per-frame guidance scale [T] (fp32) for the fused combine+Euler kernel; None if not expressible."""
        T = x.shape[0]
        s = self.scale_schedule(sigma, self.scale_rule(scale, **guider_kwargs))
        if not isinstance(s, torch.Tensor):
            return torch.full((T,), float(s), device=x.device, dtype=torch.float32)
        if s.numel() == T:
            return s.reshape(T).to(device=x.device, dtype=torch.float32).contiguous()
        if s.numel() == 1:
            return s.reshape(1).to(device=x.device, dtype=torch.float32).expand(T).contiguous()
        return None


class MultiviewCFG(VanillaCFG):
    def __init__(self, cfg_min: float = 1.0):
        self.scale_min = cfg_min
        self.scale_rule = MultiviewScaleRule(min_scale=cfg_min)
        self.scale_schedule = ConstantScaleSchedule()
        self.guidance = ConstantGuidance()

    def __call__(self, x: torch.Tensor, sigma: float | torch.Tensor, scale: float | torch.Tensor,  # type: ignore
                 c2w: torch.Tensor, K: torch.Tensor, input_frame_mask: torch.Tensor) -> torch.Tensor:
        x_u, x_c = x.chunk(2)
        scale = self.scale_rule(scale, c2w, K, input_frame_mask)
        scale_value = self.scale_schedule(sigma, scale)
        return self.guidance(x_u, x_c, scale_value)


class MultiviewTemporalCFG(MultiviewCFG):
    def __init__(self, num_frames: int, cfg_min: float = 1.0):
        super().__init__(cfg_min=cfg_min)
        self.num_frames = num_frames
        ar = torch.arange(num_frames)
        self.distance_matrix = (ar[None] - ar[:, None]).abs()

    def temporal_scale(self, scale, input_frame_mask: torch.Tensor, ndim: int):
        """Scale ramp with the index distance to the nearest input frame (sampling.py:286-296)."""
        T = self.num_frames
        mask = input_frame_mask.reshape(-1, T)
        min_distance = (self.distance_matrix[None].to(mask.device) + (~mask[:, None]) * T).min(-1)[0]
        min_distance = min_distance / min_distance.max(-1, keepdim=True)[0].clamp(min=1)
        scale = min_distance * (scale - self.scale_min) + self.scale_min
        return append_dims(scale.reshape(-1), ndim), mask.flatten(0, 1)

    def __call__(self, x: torch.Tensor, sigma: float | torch.Tensor, scale: float | torch.Tensor,
                 c2w: torch.Tensor, K: torch.Tensor, input_frame_mask: torch.Tensor) -> torch.Tensor:
        scale, mask = self.temporal_scale(scale, input_frame_mask, x.ndim)
        return super().__call__(x, sigma, scale, c2w, K, mask)

    def frame_scale(self, x, sigma, scale, c2w=None, K=None, input_frame_mask=None, **_):
        scale, mask = self.temporal_scale(scale, input_frame_mask, x.ndim)
        return super().frame_scale(x, sigma, scale, c2w=c2w, K=K, input_frame_mask=mask)


class EulerEDMSampler(object):
    def __init__(self, discretization: DDPMDiscretization, guider: VanillaCFG | MultiviewCFG | MultiviewTemporalCFG,
                 num_steps: int | None = None, verbose: bool = False, device: str | torch.device = "cuda",
                 s_churn=0.0, s_tmin=0.0, s_tmax=float("inf"), s_noise=1.0):
        self.num_steps = num_steps
        self.discretization = discretization
        self.guider = guider
        self.verbose = verbose
        self.device = device
        self.s_churn, self.s_tmin, self.s_tmax, self.s_noise = s_churn, s_tmin, s_tmax, s_noise
        self._scale_cache = None

    def prepare_sampling_loop(self, x: torch.Tensor, cond: dict, uc: dict, num_steps: int | None = None):
        num_steps = num_steps or self.num_steps
        assert num_steps is not None, "num_steps must be specified"
        sigmas = self.discretization(num_steps, device=self.device)
        x *= torch.sqrt(1.0 + sigmas[0] ** 2.0)
        num_sigmas = len(sigmas)
        s_in = x.new_ones([x.shape[0]])
        self._scale_cache = None
        reset = getattr(self.guider, "reset_cache", None)
        if reset is not None:  # conditioning may have been rewritten in place since the last loop
            reset()
        return x, s_in, sigmas, num_sigmas, cond, uc

    def get_sigma_gen(self, num_sigmas: int, verbose: bool = True):
        sigma_generator = range(num_sigmas - 1)
        if self.verbose and verbose and tqdm is not None:
            sigma_generator = tqdm(sigma_generator, total=num_sigmas - 1, desc="Sampling", leave=False)
        return sigma_generator

    def _fused_scale(self, x, sigma_hat, scale, guider_kwargs):
        """Per-frame guidance scale for the fused kernel; it does not depend on the step, so it is computed once
        per sampling run (cache dropped in prepare_sampling_loop)."""
        if not hasattr(self.guider, "frame_scale"):
            return None
        key = (id(scale), tuple(sorted((k, id(v)) for k, v in guider_kwargs.items())), x.shape[0])
        if self._scale_cache is not None and self._scale_cache[0] == key:
            return self._scale_cache[1]
        try:
            s = self.guider.frame_scale(x, sigma_hat, scale, **guider_kwargs)
        except TypeError:
            s = None
        self._scale_cache = (key, s, scale, guider_kwargs)  # hold refs so ids stay unique
        return s

    def sampler_step(self, sigma: torch.Tensor, next_sigma: torch.Tensor, denoiser, x: torch.Tensor,
                     scale: float | torch.Tensor, cond: dict, uc: dict, gamma: float = 0.0,
                     **guider_kwargs) -> torch.Tensor:
        sigma_hat = sigma * (gamma + 1.0) + 1e-6
        eps = torch.randn_like(x) * self.s_noise
        fused = x.is_cuda and x.dtype == torch.float32 and x.ndim == 4
        if fused:
            x = ops.add_noise(x, eps, sigma, sigma_hat)
        else:
            x = x + eps * append_dims(sigma_hat**2 - sigma**2, x.ndim) ** 0.5

        denoised = denoiser(*self.guider.prepare_inputs(x, sigma_hat, cond, uc))
        guide = self.guider
        if hasattr(guide, "gather"):  # dist.CFGSplit: this rank denoised one half; restore [uncond; cond]
            denoised, guide = guide.gather(denoised), guide.guider
        frame_scale = self._fused_scale(x, sigma_hat, scale, guider_kwargs) if fused else None
        if frame_scale is not None and denoised.dtype == torch.float32 and denoised.shape[0] == 2 * x.shape[0] \
                and (x[0].numel() % 4) == 0:
            return ops.cfg_euler_step(x, denoised, frame_scale, sigma_hat.contiguous(), next_sigma.contiguous())
        denoised = guide(denoised, sigma_hat, scale, **guider_kwargs)
        d = to_d(x, sigma_hat, denoised)
        dt = append_dims(next_sigma - sigma_hat, x.ndim)
        return x + dt * d

    def __call__(self, denoiser, x: torch.Tensor, scale: float | torch.Tensor, cond: dict, uc: dict | None = None,
                 num_steps: int | None = None, verbose: bool = True, **guider_kwargs) -> torch.Tensor:
        uc = cond if uc is None else uc
        x, s_in, sigmas, num_sigmas, cond, uc = self.prepare_sampling_loop(x, cond, uc, num_steps)
        sig_host = sigmas.detach().cpu().tolist()  # one transfer instead of a tensor->bool sync per step
        for i in self.get_sigma_gen(num_sigmas, verbose=verbose):
            gamma = (min(self.s_churn / (num_sigmas - 1), 2**0.5 - 1)
                     if self.s_tmin <= sig_host[i] <= self.s_tmax else 0.0)
            x = self.sampler_step(s_in * sigmas[i], s_in * sigmas[i + 1], denoiser, x, scale, cond, uc, gamma,
                                  **guider_kwargs)
        return x
