"""ORACLE (test infrastructure only — never imported by the product path).

CPU restatement (numpy for the schedules, torch fp32 for the tensor updates) of the reference's sampler loops:
  * schedule buffers  — ldm/models/diffusion/ddpm.py:111-161, ldm/modules/diffusionmodules/util.py:134-185
  * PLMSSampler       — ldm/models/diffusion/plms.py (whole file)
  * DPMSolverSampler  — ldm/models/diffusion/dpm_solver/sampler.py:31-91 and dpm_solver.py
                        (NoiseScheduleVP discrete 92-168, model_wrapper 171-329, DPM_Solver 332-384, 488-520,
                        742-777, 958-1123, interpolate_fn 1126-1171)
  * "DDIM"            — the reference has no DDIMSampler class; DDIM here is the update embedded
                        in PLMSSampler.get_x_prev_and_pred_x0 (plms.py:210-228) applied with e' = e_t each step.
`model_fn(x, t, cond)` stands for `model.apply_model` (ddpm.py:290-306): x [n,C,H,W], t [n], cond [n,L,D] -> eps.

Pinning: see oracle/sd_unet.py; the same golden-vector script runs the unmodified reference samplers.

Precision notes: the reference stores the schedule in the model dtype and, in the DPM-Solver path,
rounds the time grid and `t_array` to fp16 unconditionally (dpm_solver.py:100-103,415) and computes the model
input time in fp16 (dpm_solver.py:256-265 on an fp16 tensor), and rounds x_T to fp16 once (sampler.py:89; the state
is promoted back to fp32 by the first update because eps and the coefficients are fp32).  All of these are reproduced.
"""
import numpy as np
import torch


# ----------------------------------------------------------------------------------------------------------
# schedules
# ----------------------------------------------------------------------------------------------------------
def make_beta_schedule(n_timestep=1000, linear_start=0.00085, linear_end=0.012):
    """util.py:172-185 ('linear' = linear in sqrt(beta)); LinSpace runs in fp32."""
    start = np.float32(linear_start ** 0.5)
    stop = np.float32(linear_end ** 0.5)
    lin = np.linspace(start, stop, n_timestep, dtype=np.float32)
    return (lin ** 2).astype(np.float32)


def register_schedule(n_timestep=1000, linear_start=0.00085, linear_end=0.012, dtype=np.float32):
    """ddpm.py:111-139. Returns dict with betas, alphas_cumprod, alphas_cumprod_prev in `dtype`."""
    betas = make_beta_schedule(n_timestep, linear_start, linear_end)
    alphas = 1.0 - betas
    alphas_cumprod = np.cumprod(alphas, axis=0)
    alphas_cumprod_prev = np.append(1.0, alphas_cumprod[:-1])
    return dict(betas=betas.astype(dtype), alphas_cumprod=alphas_cumprod.astype(dtype),
                alphas_cumprod_prev=alphas_cumprod_prev.astype(dtype), num_timesteps=int(betas.shape[0]))


def make_ddim_timesteps(num_ddim_timesteps, num_ddpm_timesteps=1000):
    """util.py:134-148, 'uniform'."""
    c = num_ddpm_timesteps // num_ddim_timesteps
    return np.asarray(list(range(0, num_ddpm_timesteps, c))) + 1


def make_ddim_sampling_parameters(alphacums, ddim_timesteps, eta=0.0):
    """util.py:151-162."""
    alphas = alphacums[ddim_timesteps]
    alphas_prev = np.asarray([alphacums[0]] + alphacums[ddim_timesteps[:-1]].tolist(), dtype=alphacums.dtype)
    sigmas = eta * np.sqrt((1 - alphas_prev) / (1 - alphas) * (1 - alphas / alphas_prev))
    return sigmas.astype(alphacums.dtype), alphas, alphas_prev


# ----------------------------------------------------------------------------------------------------------
# PLMS / DDIM
# ----------------------------------------------------------------------------------------------------------
def _cfg_model_output(model_fn, x, t, c, uc, scale, v_coef=None):
    """get_model_output, plms.py:188-203.  v_coef = (sqrt(abar), sqrt(1 - abar)) [1000]: the network predicts v and
    the guided output is converted with model_wrapper's rule eps = alpha_t * v + sigma_t * x (dpm_solver.py:281-284;
    the reference's PLMS itself only knows eps networks, plms.py:200)."""
    if uc is None or scale == 1.0:
        out = model_fn(x, t, c)
    else:
        x_in = torch.cat((x, x), dim=0)
        t_in = torch.cat((t, t), dim=0)
        c_in = torch.cat((uc, c), dim=0)
        e_u, e_c = model_fn(x_in, t_in, c_in).chunk(2, dim=0)
        out = e_u + scale * (e_c - e_u)
    if v_coef is not None:
        a = torch.tensor(v_coef[0][int(t[0])], dtype=torch.float32)
        sg = torch.tensor(v_coef[1][int(t[0])], dtype=torch.float32)
        out = a * out + sg * x
    return out


def plms_sample(model_fn, sched, S, x_T, cond, uncond=None, scale=1.0, eta=0.0, ddim=False, noise_fn=None,
                log_every_t=100, record=None, model_type="noise", mask=None, x0=None, blend_noise_fn=None):
    """PLMSSampler.sample -> plms_sampling -> p_sample_plms (plms.py:69-248). With ddim=True every step uses
    e' = e_t (no multistep), which is the DDIM sampler the metric names.  Returns (img, intermediates).
    mask / x0: the inpainting blend of plms.py:153-156, `img = q_sample(x0, ts, randn) * mask + (1 - mask) * img`
    (q_sample: wukong-huahua ddpm.py:180-182); blend_noise_fn(shape) supplies the draw.  x0 without a mask is
    ignored, as in the reference."""
    if eta != 0 and not ddim:
        raise ValueError("ddim_eta must be 0 for PLMS")  # plms.py:35-36
    ac = sched["alphas_cumprod"]
    assert ac.shape[0] == sched["num_timesteps"]
    ts = make_ddim_timesteps(S, sched["num_timesteps"])
    sigmas, alphas, alphas_prev = make_ddim_sampling_parameters(ac, ts, eta)
    sqrt_1m = np.sqrt(1.0 - alphas)
    img = x_T.clone()
    b = img.shape[0]
    v_coef = (np.sqrt(ac), np.sqrt(1.0 - ac)) if model_type == "v" else None
    time_range = np.flip(ts)
    total = ts.shape[0]
    inter = {"x_inter": [img], "pred_x0": [img]}
    old_eps = []

    def x_prev_and_x0(x, e, index):  # plms.py:210-228
        a_t = torch.tensor(alphas[index], dtype=torch.float32)
        a_prev = torch.tensor(alphas_prev[index], dtype=torch.float32)
        sigma_t = torch.tensor(sigmas[index], dtype=torch.float32)
        s1m = torch.tensor(sqrt_1m[index], dtype=torch.float32)
        pred_x0 = (x - s1m * e) / a_t.sqrt()
        dir_xt = (1.0 - a_prev - sigma_t ** 2).sqrt() * e
        noise = sigma_t * (noise_fn(x.shape) if (noise_fn is not None and float(sigma_t) != 0.0) else 0.0)
        return a_prev.sqrt() * pred_x0 + dir_xt + noise, pred_x0

    for i, step in enumerate(time_range):
        index = total - i - 1
        t = torch.full((b,), int(step), dtype=torch.int64)
        t_next = torch.full((b,), int(time_range[min(i + 1, len(time_range) - 1)]), dtype=torch.int64)
        if mask is not None:
            assert x0 is not None
            a_s = torch.tensor(np.sqrt(ac)[int(step)], dtype=torch.float32)
            s_s = torch.tensor(np.sqrt(1.0 - ac)[int(step)], dtype=torch.float32)
            img_orig = a_s * x0 + s_s * blend_noise_fn(x0.shape)
            img = img_orig * mask + (1.0 - mask) * img
        e_t = _cfg_model_output(model_fn, img, t, cond, uncond, scale, v_coef)
        if record is not None:
            record.append(e_t)
        if ddim:
            e_prime = e_t
        elif len(old_eps) == 0:  # pseudo improved Euler, plms.py:231-235
            x_p, _ = x_prev_and_x0(img, e_t, index)
            e_next = _cfg_model_output(model_fn, x_p, t_next, cond, uncond, scale, v_coef)
            if record is not None:
                record.append(e_next)
            e_prime = (e_t + e_next) / 2
        elif len(old_eps) == 1:
            e_prime = (3 * e_t - old_eps[-1]) / 2
        elif len(old_eps) == 2:
            e_prime = (23 * e_t - 16 * old_eps[-1] + 5 * old_eps[-2]) / 12
        else:
            e_prime = (55 * e_t - 59 * old_eps[-1] + 37 * old_eps[-2] - 9 * old_eps[-3]) / 24
        img, pred_x0 = x_prev_and_x0(img, e_prime, index)
        old_eps.append(e_t)
        if len(old_eps) >= 4:
            old_eps.pop(0)
        if index % log_every_t == 0 or index == total - 1:
            inter["x_inter"].append(img)
            inter["pred_x0"].append(pred_x0)
    return img, inter


# ----------------------------------------------------------------------------------------------------------
# DPM-Solver++ (2M)
# ----------------------------------------------------------------------------------------------------------
def interpolate_fn(x, xp, yp):
    """dpm_solver.py:1126-1171 for C = 1: piecewise linear through (xp, yp), the outermost segments extended
    beyond the key points.  x: [N], xp/yp: [K] (xp ascending). float32 arithmetic."""
    x = np.asarray(x, dtype=np.float32)
    xp = np.asarray(xp, dtype=np.float32)
    yp = np.asarray(yp, dtype=np.float32)
    K = xp.shape[0]
    out = np.empty_like(x)
    for i, xv in enumerate(x):
        # x_idx = position of x in sort(concat(x, xp)) (ties give the same value either way)
        x_idx = int(np.searchsorted(xp, xv, side="left"))
        if x_idx == 0:
            s = 0
        elif x_idx == K:
            s = K - 2
        else:
            s = x_idx - 1
        out[i] = yp[s] + (xv - xp[s]) * (yp[s + 1] - yp[s]) / (xp[s + 1] - xp[s])
    return out


class NoiseScheduleVP:
    """dpm_solver.py:20-168, schedule='discrete'."""

    def __init__(self, alphas_cumprod):
        log_alphas = (0.5 * np.log(np.asarray(alphas_cumprod, dtype=np.float32))).astype(np.float32)
        self.total_N = len(log_alphas)
        self.T = 1.0
        t32 = np.linspace(np.float32(0.0), np.float32(1.0), self.total_N + 1, dtype=np.float32)[1:]
        self.t_array = t32.astype(np.float16).astype(np.float32)  # cast(..., ms.float16), dpm_solver.py:100-103
        self.log_alpha_array = log_alphas

    def marginal_log_mean_coeff(self, t):
        return interpolate_fn(np.atleast_1d(t), self.t_array, self.log_alpha_array)

    def marginal_alpha(self, t):
        return np.exp(self.marginal_log_mean_coeff(t))

    def marginal_std(self, t):
        return np.sqrt(1.0 - np.exp(2.0 * self.marginal_log_mean_coeff(t)))

    def marginal_lambda(self, t):
        lmc = self.marginal_log_mean_coeff(t)
        return lmc - 0.5 * np.log(1.0 - np.exp(2.0 * lmc))


def dpm_time_steps(S, t_T=1.0, t_0=1e-3):
    """get_time_steps('time_uniform'), dpm_solver.py:395-415: fp32 linspace rounded to fp16."""
    return np.linspace(np.float32(t_T), np.float32(t_0), S + 1, dtype=np.float32).astype(np.float16)


def dpm_model_input_time(t16, total_N=1000):
    """get_model_input_time on an fp16 tensor, dpm_solver.py:256-265: (t - 1/N) * 1000 evaluated in fp16."""
    t16 = np.float16(t16)
    return np.float32((t16 - np.float16(1.0 / total_N)) * np.float16(1000.0))


def dpmpp2m_sample(model_fn, sched, S, x_T, cond, uncond=None, scale=1.0, model_type="noise", record=None):
    """DPMSolverSampler.sample (sampler.py:31-91) -> DPM_Solver.sample(method='multistep', order=2,
    skip_type='time_uniform', lower_order_final=True) (dpm_solver.py:958-1094), predict_x0=True."""
    ns = NoiseScheduleVP(sched["alphas_cumprod"])
    t_0 = 1.0 / ns.total_N
    timesteps = dpm_time_steps(S, ns.T, t_0)
    x = x_T.to(torch.float16).to(torch.float32)  # ops.Cast()(img, ms.float16), sampler.py:89 (state is fp32 afterwards)
    b = x.shape[0]

    def model_eval(x, t16):  # model_wrapper.model_fn + data_prediction_fn, dpm_solver.py:302-325,366-373
        t_in = torch.full((b,), float(dpm_model_input_time(t16, ns.total_N)), dtype=torch.float32)
        if uncond is None or scale == 1.0:
            out = model_fn(x, t_in, cond)
        else:
            o_u, o_c = model_fn(torch.cat([x] * 2), torch.cat([t_in] * 2), torch.cat([uncond, cond])).chunk(2)
            out = o_u + scale * (o_c - o_u)
        tt = np.float32(t16)
        alpha_t = float(ns.marginal_alpha(tt)[0])
        sigma_t = float(ns.marginal_std(tt)[0])
        if record is not None:
            record.append(out)
        if model_type == "v":  # dpm_solver.py:281-284
            noise = alpha_t * out + sigma_t * x
        else:
            noise = out
        return (x - sigma_t * noise) / alpha_t

    def first_update(x, s, t, model_s):  # dpm_solver.py:488-520, predict_x0 branch
        lam_s, lam_t = ns.marginal_lambda(s)[0], ns.marginal_lambda(t)[0]
        h = lam_t - lam_s
        sigma_s, sigma_t = ns.marginal_std(s)[0], ns.marginal_std(t)[0]
        alpha_t = np.exp(ns.marginal_log_mean_coeff(t)[0])
        phi_1 = np.expm1(-h)
        return float(sigma_t / sigma_s) * x - float(alpha_t * phi_1) * model_s

    def second_update(x, m_prev_1, m_prev_0, t_prev_1, t_prev_0, t):  # dpm_solver.py:742-777, 'dpm_solver'
        lam_p1, lam_p0, lam_t = (ns.marginal_lambda(t_prev_1)[0], ns.marginal_lambda(t_prev_0)[0],
                                 ns.marginal_lambda(t)[0])
        sigma_p0, sigma_t = ns.marginal_std(t_prev_0)[0], ns.marginal_std(t)[0]
        alpha_t = np.exp(ns.marginal_log_mean_coeff(t)[0])
        h_0 = lam_p0 - lam_p1
        h = lam_t - lam_p0
        r0 = h_0 / h
        D1_0 = float(1.0 / r0) * (m_prev_0 - m_prev_1)
        c = float(alpha_t * (np.exp(-h) - 1.0))
        return float(sigma_t / sigma_p0) * x - c * m_prev_0 - 0.5 * c * D1_0

    order = 2
    assert S >= order
    t_prev = [np.float32(timesteps[0])]
    m_prev = [model_eval(x, timesteps[0])]
    # init: order-1 step
    x = first_update(x, t_prev[-1], np.float32(timesteps[1]), m_prev[-1])
    m_prev.append(model_eval(x, timesteps[1]))
    t_prev.append(np.float32(timesteps[1]))
    for step in range(order, S + 1):
        t = np.float32(timesteps[step])
        step_order = min(order, S + 1 - step) if S < 15 else order  # lower_order_final, dpm_solver.py:1083-1086
        if step_order == 1:
            x = first_update(x, t_prev[-1], t, m_prev[-1])
        else:
            x = second_update(x, m_prev[0], m_prev[1], t_prev[0], t_prev[1], t)
        t_prev[0], m_prev[0] = t_prev[1], m_prev[1]
        t_prev[1] = t
        if step < S:
            m_prev[1] = model_eval(x, timesteps[step])
    return x
