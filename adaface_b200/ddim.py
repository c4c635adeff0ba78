"""Host-side mirror of the reference's DDIM sampler (reference: ldm/models/diffusion/ddim.py - DDIMSampler :12,
make_schedule :28-68, sample :71-133, ddim_sampling :135-220, p_sample_ddim :222-296, decode :315-349; schedule
helpers ldm/modules/diffusionmodules/util.py:46-77).

Same method names and arguments.  The classifier-free-guidance combine and the DDIM update - about ten
elementwise torch kernels and four `torch.full` allocations per step in the reference - are ONE vectorised
fp32 kernel here (ops.cfg_ddim_step, bit-identical to the reference's op order).  The schedule is host-side
scalar math; no tensor is moved to "cuda" behind the caller's back (the reference's register_buffer does,
SURVEY.md F6) - buffers stay on the model's device.
"""
from __future__ import annotations

from typing import Optional, Sequence

import numpy as np
import torch

from . import ops

Tensor = torch.Tensor


def make_ddim_timesteps(ddim_discr_method: str, num_ddim_timesteps: int, num_ddpm_timesteps: int,
                        verbose: bool = False) -> np.ndarray:
    """diffusionmodules/util.py:46-62."""
    if ddim_discr_method == "uniform":
        c = num_ddpm_timesteps // num_ddim_timesteps
        steps = np.asarray(list(range(0, num_ddpm_timesteps, c)))
    elif ddim_discr_method == "quad":
        steps = ((np.linspace(0, np.sqrt(num_ddpm_timesteps * .8), num_ddim_timesteps)) ** 2).astype(int)
    else:
        raise NotImplementedError(f'There is no ddim discretization method called "{ddim_discr_method}"')
    steps_out = steps + 1
    if verbose:
        print(f"Selected timesteps for ddim sampler: {steps_out}")
    return steps_out


def make_ddim_sampling_parameters(alphacums: np.ndarray, ddim_timesteps: np.ndarray, eta: float):
    """diffusionmodules/util.py:65-77 on float32-valued inputs; sigma is evaluated in float64 and rounded to
    float32, as the reference's mixed ndarray/tensor expression does."""
    alphacums = np.asarray(alphacums, dtype=np.float32)
    alphas = alphacums[ddim_timesteps]
    alphas_prev = np.asarray([alphacums[0]] + alphacums[ddim_timesteps[:-1]].tolist(), dtype=np.float32)
    a64, ap64 = alphas.astype(np.float64), alphas_prev.astype(np.float64)
    sigmas = (eta * np.sqrt((1 - ap64) / (1 - a64) * (1 - a64 / ap64))).astype(np.float32)
    return sigmas, alphas, alphas_prev


def guidance_schedule(guidance_scale, total_steps: int):
    """Guidance annealing of ddim_sampling (ddim.py:169-180, 215-218): the scale used at loop step i.
    A scalar guidance_scale crashes in the reference (undefined max_guide_scale); here it
    means (g, g), i.e. no annealing."""
    if isinstance(guidance_scale, (list, tuple)):
        g_max, g_min = guidance_scale
    else:
        g_max = g_min = guidance_scale
    delta = (g_max - g_min) / max(total_steps - 1, 1)
    out, g = [], g_max
    for _ in range(total_steps):
        out.append(g)
        g = g - delta
    return out


class DDIMSampler(object):
    def __init__(self, model, schedule: str = "linear", **kwargs):
        super().__init__()
        self.model = model
        self.ddpm_num_timesteps = model.num_timesteps
        self.schedule = schedule
        # the reference draws (and, at eta = 0, discards) a noise tensor every step (ddim.py:286), consuming
        # RNG state; keep that stream position identical by default
        self.match_reference_rng = True

    def register_buffer(self, name, attr):
        setattr(self, name, attr)

    def make_schedule(self, ddim_num_steps: int, ddim_discretize: str = "uniform", ddim_eta: float = 0.,
                      verbose: bool = False):
        self.ddim_timesteps = make_ddim_timesteps(ddim_discretize, ddim_num_steps, self.ddpm_num_timesteps, verbose)
        alphas_cumprod = self.model.alphas_cumprod
        if alphas_cumprod.shape[0] != self.ddpm_num_timesteps:
            raise ValueError("alphas have to be defined for each timestep")
        ac = alphas_cumprod.detach().to(torch.float32).cpu().numpy()
        to_dev = lambda a: torch.as_tensor(np.asarray(a), dtype=torch.float32, device=self.model.device)
        self.register_buffer("betas", self.model.betas.detach().to(torch.float32))
        self.register_buffer("alphas_cumprod", alphas_cumprod.detach().to(torch.float32))
        self.register_buffer("alphas_cumprod_prev", self.model.alphas_cumprod_prev.detach().to(torch.float32))
        self.register_buffer("sqrt_alphas_cumprod", to_dev(np.sqrt(ac)))
        self.register_buffer("sqrt_one_minus_alphas_cumprod", to_dev(np.sqrt(1. - ac)))
        sigmas, alphas, alphas_prev = make_ddim_sampling_parameters(ac, self.ddim_timesteps, ddim_eta)
        # host-side float32 arrays: indexed per step and passed to the kernel as scalars
        self.register_buffer("ddim_sigmas", sigmas)
        self.register_buffer("ddim_alphas", alphas)
        self.register_buffer("ddim_alphas_prev", alphas_prev)
        self.register_buffer("ddim_sqrt_one_minus_alphas", np.sqrt(1. - alphas).astype(np.float32))
        acp = self.model.alphas_cumprod_prev.detach().to(torch.float32).cpu().numpy().astype(np.float64)
        a64 = ac.astype(np.float64)
        with np.errstate(divide="ignore", invalid="ignore"):
            sig_orig = ddim_eta * np.sqrt((1 - acp) / (1 - a64) * (1 - a64 / acp))
        self.register_buffer("ddim_sigmas_for_original_num_steps", sig_orig.astype(np.float32))

    @torch.no_grad()
    def sample(self, S, batch_size, shape, conditioning=None, callback=None, normals_sequence=None, img_callback=None,
               quantize_x0=False, eta=0., mask=None, x0=None, temperature=1., noise_dropout=0., score_corrector=None,
               corrector_kwargs=None, verbose=False, x_T=None, log_every_t=100, guidance_scale=1.,
               unconditional_conditioning=None, **kwargs):
        self.make_schedule(ddim_num_steps=S, ddim_eta=eta, verbose=verbose)
        C, H, W = shape
        size = (batch_size, C, H, W)
        return self.ddim_sampling(conditioning, size, callback=callback, img_callback=img_callback,
                                  quantize_denoised=quantize_x0, mask=mask, x0=x0, ddim_use_original_steps=False,
                                  noise_dropout=noise_dropout, temperature=temperature, score_corrector=score_corrector,
                                  corrector_kwargs=corrector_kwargs, x_T=x_T, log_every_t=log_every_t,
                                  guidance_scale=guidance_scale, unconditional_conditioning=unconditional_conditioning,
                                  **kwargs)

    @torch.no_grad()
    def ddim_sampling(self, cond, shape, x_T=None, ddim_use_original_steps=False, callback=None, timesteps=None,
                      quantize_denoised=False, mask=None, x0=None, img_callback=None, log_every_t=100, temperature=1.,
                      noise_dropout=0., score_corrector=None, corrector_kwargs=None, guidance_scale=1.,
                      unconditional_conditioning=None, **kwargs):
        device = self.model.betas.device
        b = shape[0]
        img = torch.randn(shape, device=device) if x_T is None else x_T
        if timesteps is None:
            timesteps = self.ddpm_num_timesteps if ddim_use_original_steps else self.ddim_timesteps
        elif not ddim_use_original_steps:
            subset_end = int(min(timesteps / self.ddim_timesteps.shape[0], 1) * self.ddim_timesteps.shape[0]) - 1
            timesteps = self.ddim_timesteps[:subset_end]
        intermediates = {"x_inter": [img], "pred_x0": [img]}
        time_range = list(reversed(range(0, timesteps))) if ddim_use_original_steps else np.flip(timesteps)
        total_steps = timesteps if ddim_use_original_steps else timesteps.shape[0]
        scales = guidance_schedule(guidance_scale, total_steps)
        for i, step in enumerate(time_range):
            index = total_steps - i - 1
            ts = torch.full((b,), int(step), device=device, dtype=torch.long)
            if mask is not None:                                                                 # ddim.py:189-194
                if x0 is None:
                    raise ValueError("ddim_sampling: mask needs x0")
                img_orig = self.model.q_sample(x0, ts)
                img = img_orig * mask + (1. - mask) * img
            img, pred_x0 = self.p_sample_ddim(img, cond, ts, index=index, use_original_steps=ddim_use_original_steps,
                                              quantize_denoised=quantize_denoised, temperature=temperature,
                                              noise_dropout=noise_dropout, score_corrector=score_corrector,
                                              corrector_kwargs=corrector_kwargs, guidance_scale=scales[i],
                                              unconditional_conditioning=unconditional_conditioning, **kwargs)
            if callback:
                callback(i)
            if img_callback:
                img_callback(pred_x0, i)
            if index % log_every_t == 0 or index == total_steps - 1:
                intermediates["x_inter"].append(img)
                intermediates["pred_x0"].append(pred_x0)
        return img, intermediates

    def step_scalars(self, index: int, use_original_steps: bool = False):
        """(a_t, a_prev, sigma_t, sqrt(1 - a_t)) as python floats holding float32 values (ddim.py:267-276)."""
        if use_original_steps:
            f = lambda t: float(t[index].float().cpu()) if isinstance(t, torch.Tensor) else float(np.float32(t[index]))
            return (f(self.model.alphas_cumprod), f(self.model.alphas_cumprod_prev),
                    f(self.ddim_sigmas_for_original_num_steps), f(self.sqrt_one_minus_alphas_cumprod))
        return (float(self.ddim_alphas[index]), float(self.ddim_alphas_prev[index]), float(self.ddim_sigmas[index]),
                float(self.ddim_sqrt_one_minus_alphas[index]))

    @torch.no_grad()
    def _cat_conditioning(self, c_c, c_u):
        """torch.cat([cond, uncond]) (ddim.py:231-242), kept across the steps of a run: the prompts do not change between
        DDIM steps, and 2 x 30 MB of layerwise context per step is copying for nothing.  Keyed by the tensors' identity
        and version counters (an in-place update, e.g. a fresh host-to-device copy, concatenates again); returning the
        SAME tensor also lets DiffusionModel skip the copy into its graph's input buffer."""
        if torch.is_grad_enabled() and (c_c.requires_grad or c_u.requires_grad):
            return torch.cat([c_c, c_u])
        key = (id(c_c), c_c._version, id(c_u), c_u._version)
        cached = self.__dict__.get("_cat_cache")
        if cached is None or cached[0] != key:
            cached = (key, torch.cat([c_c, c_u]), c_c, c_u)      # the references keep the ids from being reused
            self.__dict__["_cat_cache"] = cached
        return cached[1]

    def p_sample_ddim(self, x, c, t, index, repeat_noise=False, use_original_steps=False, quantize_denoised=False,
                      temperature=1., noise_dropout=0., score_corrector=None, corrector_kwargs=None, guidance_scale=1.,
                      unconditional_conditioning=None):
        if score_corrector is not None or quantize_denoised:
            raise NotImplementedError("adaface_b200.DDIMSampler: score_corrector / quantize_denoised need the first-stage "
                                      "model, which is outside this path (unused by scripts/stable_txt2img.py)")
        if unconditional_conditioning is None or guidance_scale == 1.:
            e_c = e_u = self.model.apply_model(x, t, c)                                          # ddim.py:228
            scale = 1.0
        else:
            x_in = torch.cat([x] * 2)                                                            # ddim.py:231-232
            t_in = torch.cat([t] * 2)
            if isinstance(c, tuple):                                                             # (cond, uncond) order: F8
                c_c, c_in_c, extra_info = c
                c_u, c_in_u, _ = unconditional_conditioning
                c2 = (self._cat_conditioning(c_c, c_u), sum([list(c_in_c), list(c_in_u)], []), extra_info)
            else:
                c2 = self._cat_conditioning(c, unconditional_conditioning)
            e_c, e_u = self.model.apply_model(x_in, t_in, c2).chunk(2)                           # ddim.py:252
            scale = float(guidance_scale)
        a_t, a_prev, sigma_t, sqrt_one_minus_at = self.step_scalars(index, use_original_steps)
        noise = None
        if sigma_t != 0. or self.match_reference_rng:
            if repeat_noise:
                noise = torch.randn((1, *x.shape[1:]), device=x.device).repeat(x.shape[0], *((1,) * (x.dim() - 1)))
            else:
                noise = torch.randn(x.shape, device=x.device)
            if noise_dropout > 0.:
                noise = torch.nn.functional.dropout(noise, p=noise_dropout)
            if sigma_t == 0.:
                noise = None
        x_prev, pred_x0 = ops.cfg_ddim_step(x.float(), e_c.float(), e_u.float(), noise, scale, a_t, a_prev, sigma_t,
                                            sqrt_one_minus_at, float(temperature))               # ddim.py:260-295
        return x_prev, pred_x0

    @torch.no_grad()
    def stochastic_encode(self, x0, t, use_original_steps=False, noise=None):
        """ddim.py:298-313 (edge utility; plain tensor arithmetic on the caller's device)."""
        if use_original_steps:
            sa, s1m = self.sqrt_alphas_cumprod, self.sqrt_one_minus_alphas_cumprod
        else:
            sa = torch.as_tensor(np.sqrt(self.ddim_alphas), device=x0.device)
            s1m = torch.as_tensor(self.ddim_sqrt_one_minus_alphas, device=x0.device)
        if noise is None:
            noise = torch.randn_like(x0)
        shape = (x0.shape[0],) + (1,) * (x0.dim() - 1)
        return sa.to(x0.device)[t].reshape(shape) * x0 + s1m.to(x0.device)[t].reshape(shape) * noise

    @torch.no_grad()
    def decode(self, x_latent, cond, t_start, guidance_scale=1.0, unconditional_conditioning=None,
               use_original_steps=False):
        """ddim.py:315-349: anneals from guidance_scale down to min(2, guidance_scale)."""
        timesteps = np.arange(self.ddpm_num_timesteps) if use_original_steps else self.ddim_timesteps
        timesteps = timesteps[:t_start]
        total_steps = timesteps.shape[0]
        scales = guidance_schedule((guidance_scale, min(2.0, guidance_scale)), total_steps)
        x_dec = x_latent
        for i, step in enumerate(np.flip(timesteps)):
            index = total_steps - i - 1
            ts = torch.full((x_latent.shape[0],), int(step), device=x_latent.device, dtype=torch.long)
            x_dec, _ = self.p_sample_ddim(x_dec, cond, ts, index=index, use_original_steps=use_original_steps,
                                          guidance_scale=scales[i], unconditional_conditioning=unconditional_conditioning)
        return x_dec
