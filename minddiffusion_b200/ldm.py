"""LatentDiffusion for the sampling path: schedule buffers + apply_model, as the samplers consume them.

Mirror of the inference-time surface of the reference's `DDPM` / `LatentDiffusion`
(vision/stablediffusionv2/ldm/models/diffusion/ddpm.py:34-161, 274-306; wukong-huahua ddpm.py:276-278 for the
keyword form of apply_model).  Training (`p_losses`, EMA, optimizers) is out of scope.
"""
import numpy as np
import torch

from . import ops


def make_beta_schedule(schedule="linear", n_timestep=1000, linear_start=1e-4, linear_end=2e-2, cosine_s=8e-3):
    """ldm/modules/diffusionmodules/util.py:172-185 — 'linear' is linear in sqrt(beta), evaluated in fp32."""
    if schedule != "linear":
        raise ValueError("schedule '%s' unknown." % schedule)
    lin = np.linspace(np.float32(linear_start ** 0.5), np.float32(linear_end ** 0.5), n_timestep, dtype=np.float32)
    return (lin ** 2).astype(np.float32)


class LatentDiffusion:
    def __init__(self, unet, timesteps=1000, beta_schedule="linear", linear_start=1e-4, linear_end=2e-2,
                 cosine_s=8e-3, given_betas=None, parameterization="eps", conditioning_key="crossattn",
                 scale_factor=1.0, first_stage_model=None, cond_stage_model=None, channels=4, image_size=64,
                 v_posterior=0.0):
        # ddpm.py:65 asserts ["eps", "x0"]; "v" is the extension BASELINE config #4 asks for,
        # its arithmetic follows model_wrapper's "v" branch (dpm_solver.py:281-284).
        if parameterization not in ("eps", "x0", "v"):
            raise AssertionError('currently only supporting "eps", "x0" and "v"')
        if conditioning_key not in (None, "concat", "crossattn", "hybrid", "adm"):
            raise AssertionError("bad conditioning_key")  # ddpm.py:351
        self.model = unet
        self.unet = unet
        self.parameterization = parameterization
        self.conditioning_key = conditioning_key
        self.scale_factor = scale_factor
        self.first_stage_model = first_stage_model
        self.cond_stage_model = cond_stage_model
        self.channels = channels
        self.image_size = image_size
        self.v_posterior = v_posterior
        self.register_schedule(given_betas, beta_schedule, timesteps, linear_start, linear_end, cosine_s)

    def register_schedule(self, given_betas=None, beta_schedule="linear", timesteps=1000, linear_start=1e-4,
                          linear_end=2e-2, cosine_s=8e-3):
        """ddpm.py:111-161 (buffers the samplers read; fp32 = the reference with use_fp16: False)."""
        betas = np.asarray(given_betas, dtype=np.float32) if given_betas is not None else \
            make_beta_schedule(beta_schedule, timesteps, linear_start, linear_end, cosine_s)
        alphas = 1.0 - betas
        alphas_cumprod = np.cumprod(alphas, axis=0)
        alphas_cumprod_prev = np.append(1.0, alphas_cumprod[:-1])
        self.num_timesteps = int(betas.shape[0])
        self.linear_start, self.linear_end = linear_start, linear_end
        assert alphas_cumprod.shape[0] == self.num_timesteps, "alphas have to be defined for each timestep"
        f = lambda a: np.asarray(a, dtype=np.float32)
        self.betas = f(betas)
        self.alphas_cumprod = f(alphas_cumprod)
        self.alphas_cumprod_prev = f(alphas_cumprod_prev)
        self.sqrt_alphas_cumprod = f(np.sqrt(alphas_cumprod))
        self.sqrt_one_minus_alphas_cumprod = f(np.sqrt(1.0 - alphas_cumprod))
        self.log_one_minus_alphas_cumprod = f(np.log(1.0 - alphas_cumprod))
        self.sqrt_recip_alphas_cumprod = f(np.sqrt(1.0 / alphas_cumprod))
        self.sqrt_recipm1_alphas_cumprod = f(np.sqrt(1.0 / alphas_cumprod - 1))
        posterior_variance = (1 - self.v_posterior) * betas * (1.0 - alphas_cumprod_prev) / (1.0 - alphas_cumprod) \
            + self.v_posterior * betas
        self.posterior_variance = f(posterior_variance)
        self.posterior_log_variance_clipped = f(np.log(np.maximum(posterior_variance, 1e-20)))
        self.posterior_mean_coef1 = f(betas * np.sqrt(alphas_cumprod_prev) / (1.0 - alphas_cumprod))
        self.posterior_mean_coef2 = f((1.0 - alphas_cumprod_prev) * np.sqrt(alphas) / (1.0 - alphas_cumprod))

    # ---- reference-facing calls ----
    def apply_model(self, x_noisy, t, cond=None, c_concat=None, c_crossattn=None, return_ids=False):
        """ddpm.py:290-306 (SD2 positional cond) / wukong ddpm.py:276-278 (keywords). NCHW in, NCHW eps out."""
        if isinstance(cond, dict):
            c_concat = cond.get("c_concat", c_concat)
            c_crossattn = cond.get("c_crossattn", c_crossattn)
        elif cond is not None:
            if self.conditioning_key == "concat":
                c_concat = cond
            else:
                c_crossattn = cond
        if isinstance(c_crossattn, (list, tuple)):
            c_crossattn = c_crossattn[0]
        if isinstance(c_concat, (list, tuple)):
            c_concat = c_concat[0]
        if c_concat is None:
            return self.unet(x_noisy, t, context=c_crossattn)
        self.unet.set_concat(c_concat)  # DiffusionWrapper 'concat'/'hybrid' (wukong ddpm.py:364-372), no torch.cat
        try:
            return self.unet(x_noisy, t, context=c_crossattn)
        finally:
            self.unet.set_concat(None)

    def get_learned_conditioning(self, c):
        if self.cond_stage_model is None:
            raise RuntimeError("no cond_stage_model attached (text tower); pass precomputed conditioning")
        return self.cond_stage_model.encode(c)

    def decode_first_stage(self, z):
        if self.first_stage_model is None:
            raise RuntimeError("no first_stage_model attached (VAE decoder)")
        return self.first_stage_model.decode(1.0 / self.scale_factor * z)
