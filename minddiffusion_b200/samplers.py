"""Sampler loops with the reference's API, running the UNet and the per-step arithmetic on the GPU kernels.

    sampler = PLMSSampler(model)            # or DPMSolverSampler(model) / DDIMSampler(model)
    samples, intermediates = sampler.sample(S, batch_size, shape, conditioning, ..., x_T=...,
                                            unconditional_guidance_scale=..., unconditional_conditioning=..., eta=...)

Signatures, argument meaning, printed warnings and raised errors follow
vision/stablediffusionv2/ldm/models/diffusion/plms.py:27-248 and dpm_solver/sampler.py:21-92 (wukong-huahua's
copies are identical for this path).  `DDIMSampler` has no class in the reference: it is the DDIM
update of `PLMSSampler.get_x_prev_and_pred_x0` (plms.py:210-228) applied with e' = e_t every step, the sampler the
headline metric names.

Per step the host does: one UNet evaluation (a CUDA-graph replay when enabled) and ONE elementwise kernel
(classifier-free-guidance combine + latent update, mdx_sampler_*).  The CFG concat of plms.py:192-194 is not
materialised: the UNet reads the latent batch twice.
"""
import numpy as np
import torch

from . import _lib, ops


def make_ddim_timesteps(ddim_discr_method="uniform", num_ddim_timesteps=50, num_ddpm_timesteps=1000, verbose=False):
    """ldm/modules/diffusionmodules/util.py:134-148."""
    if ddim_discr_method == "uniform":
        c = num_ddpm_timesteps // num_ddim_timesteps
        ddim_timesteps = np.asarray(list(range(0, num_ddpm_timesteps, c)))
    elif ddim_discr_method == "quad":
        ddim_timesteps = ((np.linspace(0, np.sqrt(num_ddpm_timesteps * .8), num_ddim_timesteps)) ** 2).astype(int)
    else:
        raise NotImplementedError('There is no ddim discretization method called "%s"' % ddim_discr_method)
    steps_out = ddim_timesteps + 1
    if verbose:
        print("Selected timesteps for ddim sampler: %s" % steps_out)
    return steps_out


def make_ddim_sampling_parameters(alphacums, ddim_timesteps, eta=0.0, verbose=False):
    """ldm/modules/diffusionmodules/util.py:151-162."""
    alphacums = np.asarray(alphacums, dtype=np.float32)
    alphas = alphacums[ddim_timesteps]
    alphas_prev = np.concatenate((alphacums[:1], alphacums[ddim_timesteps[:-1]]))
    sigmas = (eta * np.sqrt((1 - alphas_prev) / (1 - alphas) * (1 - alphas / alphas_prev))).astype(np.float32)
    return sigmas, alphas, alphas_prev


def _check_unsupported(mask, x0, score_corrector, quantize_x0, noise_dropout):
    if mask is not None and x0 is None:
        raise AssertionError("x0 must be given with mask")  # plms.py:154 `assert x0 is not None`
    # x0 without a mask is ignored, as the reference does (plms.py:153; inpaint.py:95-106 calls it that way)
    if score_corrector is not None:
        raise NotImplementedError("score_corrector is not supported")
    if quantize_x0:
        raise NotImplementedError("quantize_x0 is not supported")
    if noise_dropout:
        raise NotImplementedError("noise_dropout is not supported")


def _model_type(model):
    """0: the network predicts eps; 1: it predicts v and the step kernel converts (dpm_solver.py:281-284).  The
    reference's PLMS assumes eps (plms.py:200); "x0" networks are not used by any shipped config."""
    par = getattr(model, "parameterization", "eps")
    if par == "eps":
        return 0
    if par == "v":
        return 1
    raise NotImplementedError('parameterization "%s" is not supported by the sampler kernels' % par)


def _warn_batch(conditioning, batch_size):
    if conditioning is not None:  # plms.py:93-100
        if isinstance(conditioning, dict):
            cbs = conditioning[list(conditioning.keys())[0]]
            cbs = (cbs[0] if isinstance(cbs, (list, tuple)) else cbs).shape[0]
        else:
            cbs = conditioning.shape[0]
        if cbs != batch_size:
            print("Warning: Got %d conditionings but batch-size is %d" % (cbs, batch_size))


MAX_ROWS = 1002   # table rows of a step plan: one per UNet evaluation of a loop (S <= 1000, + the pseudo-Euler extra)


class _StepPlan:
    """State and launch sequences of one sampler loop on one latent shape.

    A step is: timestep row -> UNet evaluation (the guided batch reads the latent twice) -> ONE combine + update
    kernel -> counter advance.  Every per-step scalar lives in device tables (`table`: kernel coefficients, `t_table`:
    model-input times) indexed by the device `counter`, and every tensor the step touches has a fixed address, so
    the launches of a step are the same from step to step: with `unet.use_cuda_graph` each phase is captured once and
    the host loop only replays it; without, the same functions run eagerly.

    Phases: "first" (PLMS pseudo improved Euler, plms.py:231-235: two evaluations), "multi" (PLMS Adams-Bashforth
    2/3/4 — the weights are table entries, the eps history shifts inside the kernel), "ddim", "dpm".
    """

    def __init__(self, unet, B, C, H, W, cfg, scale, model_type, split):
        dev = unet.device
        z = lambda: torch.zeros((B, C, H, W), dtype=torch.float32, device=dev)
        self.unet, self.C, self.cfg, self.scale, self.model_type, self.split = unet, C, cfg, float(scale), model_type, split
        self.img, self.x_tmp, self.pred_x0, self.noise = z(), z(), z(), z()
        self.hist = [z(), z(), z()]
        self.table = torch.zeros((MAX_ROWS, ops.STEP_COEFS), dtype=torch.float32, device=dev)
        self.t_table = torch.zeros((MAX_ROWS,), dtype=torch.float32, device=dev)
        self.counter = torch.zeros(1, dtype=torch.int32, device=dev)
        self.repeat = 2 if (cfg and split is None) else 1
        self.t = torch.zeros(B * self.repeat, dtype=torch.float32, device=dev)
        self.model_out = None
        self.emb_table = self.emb_row = None
        if split is not None:   # eps halves: this rank's (4 real channels of the padded conv_out) and the pair's
            self.packed = torch.zeros((B, H, W, C), dtype=torch.float32, device=dev)
            self.xbuf = torch.zeros((2 * B, H, W, C), dtype=torch.float32, device=dev)
        self.graphs = {}
        if unet._ctx is None or unet._ctx["n"] != B * self.repeat:
            raise ValueError("context batch does not match the latent batch")

    def load_tables(self, rows, times):
        if len(rows) > MAX_ROWS:
            raise ValueError("too many steps for a step plan")
        self.table[:len(rows)].copy_(torch.tensor(rows, dtype=torch.float32))
        self.t_table[:len(times)].copy_(torch.tensor(times, dtype=torch.float32))
        # the timestep-embedding MLPs of every evaluation of the loop, in one batch (they depend only on t)
        emb = self.unet.emb_rows(self.t_table[:len(times)])
        if self.emb_table is None or self.emb_table.shape[1] != emb.shape[1]:
            self.emb_table = torch.zeros((MAX_ROWS, emb.shape[1]), dtype=torch.float32, device=emb.device)
            self.emb_row = torch.zeros((1, emb.shape[1]), dtype=torch.float32, device=emb.device)
            self.graphs = {}
        self.emb_table[:len(times)].copy_(emb)
        self.counter.zero_()
        for h in self.hist:
            h.zero_()

    # ---- launch sequences (identical eager and under capture) ----
    # A phase is a list of segments.  Without the CFG pair split (or with split.in_graph) it is ONE segment; with it,
    # the NCCL exchange of the eps halves sits between two captured segments and is issued eagerly.
    def _eval(self, x):
        ops.gather_row(self.emb_table, self.counter, self.emb_row)     # this evaluation's embedding row
        self.model_out = self.unet._forward_impl(x, self.t, self.repeat, None, True, emb_row=self.emb_row)
        if self.split is not None:
            self.split.pack(self.model_out, self.packed, self.C)

    def _exchange(self, dry=False):
        if not dry:
            self.split.all_gather(self.packed, self.xbuf)
        self.model_out = self.xbuf

    def _kernel(self, kind, x_out=None, **kw):
        ops.sampler_step_ex(kind, self.model_out, self.img, self.img if x_out is None else x_out, self.C, self.cfg,
                            self.scale, table=self.table, counter=self.counter, model_type=self.model_type, **kw)
        ops.step_advance(self.counter, 1)

    def _phase(self, phase):
        """[(callable, is_exchange)] in launch order."""
        ev = lambda x: ((lambda: self._eval(x)), False)
        k = lambda kind, **kw: ((lambda: self._kernel(kind, **kw)), False)
        xch = [(self._exchange, True)] if self.split is not None else []
        if phase == "first":     # PLMS pseudo improved Euler: evaluate, Euler step to x_tmp, evaluate there, average
            return ([ev(self.img)] + xch + [k("plms", x_out=self.x_tmp, e_t_out=self.hist[0]), ev(self.x_tmp)] + xch +
                    [k("plms", x_model=self.x_tmp, hist=(self.hist[0], None, None), pred_x0=self.pred_x0)])
        if phase == "multi":     # Adams-Bashforth 2 / 3 / 4 (weights in the table), history shifts in the kernel
            return [ev(self.img)] + xch + [k("plms", hist=tuple(self.hist), shift_hist=True, pred_x0=self.pred_x0)]
        if phase == "ddim":
            return [ev(self.img)] + xch + [k("plms", noise=self.noise, pred_x0=self.pred_x0)]
        if phase == "dpm":       # x0_cur overwrites x0_prev in place
            return [ev(self.img)] + xch + [k("dpm", hist=(self.hist[0], None, None), pred_x0=self.hist[0])]
        raise ValueError(phase)

    def _segments(self, phase):
        """Consecutive launches merged into capturable segments; an exchange is a segment of its own unless the split
        asks for it inside the graph."""
        in_graph = self.split is not None and self.split.in_graph
        segs, cur = [], []
        for fn, is_x in self._phase(phase):
            if is_x and not in_graph:
                if cur:
                    segs.append(("launch", cur))
                segs.append(("exchange", [fn]))
                cur = []
            else:
                cur.append(fn)
        if cur:
            segs.append(("launch", cur))
        return segs

    def run(self, phase):
        if not self.unet.use_cuda_graph:
            for fn, _ in self._phase(phase):
                fn()
            return
        ent = self.graphs.get(phase)
        if ent is None:
            # warm-up (kernel attributes, allocator, NCCL) on a snapshot of the loop state, then capture segment by
            # segment; capture itself executes nothing
            state = [self.img, self.x_tmp, self.pred_x0, self.counter] + self.hist
            saved = [t.clone() for t in state]
            for fn, _ in self._phase(phase):
                fn()
            torch.cuda.synchronize()
            ent = []
            for kind, fns in self._segments(phase):
                if kind == "exchange":
                    fns[0](dry=True)       # only the bookkeeping: the next segment reads the gathered buffer
                    ent.append((None, fns[0], 0))
                    continue
                graph = torch.cuda.CUDAGraph()
                n0 = _lib.launch_count()
                with torch.cuda.graph(graph, pool=self.unet.graph_pool()):
                    for fn in fns:
                        fn()
                ent.append((graph, None, _lib.launch_count() - n0))
            self.graphs[phase] = ent
            for t, sv in zip(state, saved):
                t.copy_(sv)
        for graph, eager, nk in ent:
            if graph is None:
                eager()
            else:
                graph.replay()
                self.unet.replayed_launches += nk


class _GuidedUNet:
    """Binds (cond, uncond, scale) to the UNet — hoisted cross-attention K/V, concat conditioning, the optional CFG
    pair split — and hands out the step plan for a latent shape."""

    def __init__(self, model, cond, uncond, scale):
        self.model = model
        self.unet = model.unet
        self.cfg = not (uncond is None or scale == 1.0)  # plms.py:189
        self.scale = float(scale)
        c_cat = uc_cat = None
        if isinstance(cond, dict):  # 'hybrid' conditioning (wukong plms.py:191-201, inpaint.py:92-96)
            one = lambda v: v[0] if isinstance(v, (list, tuple)) else v
            c_cat, cond = one(cond.get("c_concat")), one(cond.get("c_crossattn"))
            if self.cfg:
                if not isinstance(uncond, dict):
                    raise AssertionError("unconditional_conditioning must be a dict when conditioning is one")
                uc_cat, uncond = one(uncond.get("c_concat")), one(uncond.get("c_crossattn"))
        # optional: the two halves of the guided batch on two GPUs (minddiffusion_b200.parallel.CfgPairSplit)
        self.split = getattr(model, "cfg_split", None) if self.cfg else None
        if self.split is not None:
            ctx = self.split.my_context(cond, uncond)
        else:
            ctx = torch.cat((uncond, cond), dim=0) if self.cfg else cond  # order of plms.py:194
        self.unet.set_context(ctx)
        if c_cat is None:
            self.unet.set_concat(None)
        elif self.split is not None:
            self.unet.set_concat(self.split.my_context(c_cat, uc_cat))
        else:
            self.unet.set_concat(torch.cat((uc_cat, c_cat), dim=0) if self.cfg else c_cat)

    def plan(self, shape):
        B, C, H, W = shape
        key = ("step", B, C, H, W, self.cfg, self.scale, _model_type(self.model), self.split is not None)
        plans = self.unet._graphs   # dropped together with the UNet's own graphs when a captured buffer moves
        if key not in plans:
            plans[key] = _StepPlan(self.unet, B, C, H, W, self.cfg, self.scale, _model_type(self.model), self.split)
        return plans[key]


class PLMSSampler(object):
    ddim = False

    def __init__(self, model, schedule="linear", **kwargs):
        self.model = model
        self.ddpm_num_timesteps = model.num_timesteps
        self.schedule = schedule

    def make_schedule(self, ddim_num_steps, ddim_discretize="uniform", ddim_eta=0., verbose=True):
        if ddim_eta != 0 and not self.ddim:
            raise ValueError('ddim_eta must be 0 for PLMS')
        self.ddim_timesteps = make_ddim_timesteps(ddim_discr_method=ddim_discretize,
                                                  num_ddim_timesteps=ddim_num_steps,
                                                  num_ddpm_timesteps=self.ddpm_num_timesteps, verbose=verbose)
        alphas_cumprod = np.asarray(self.model.alphas_cumprod, dtype=np.float32)
        assert alphas_cumprod.shape[0] == self.ddpm_num_timesteps, 'alphas have to be defined for each timestep'
        self.betas = self.model.betas
        self.alphas_cumprod = alphas_cumprod
        self.alphas_cumprod_prev = self.model.alphas_cumprod_prev
        self.sqrt_alphas_cumprod = np.sqrt(alphas_cumprod)
        self.sqrt_one_minus_alphas_cumprod = np.sqrt(1. - alphas_cumprod)
        self.log_one_minus_alphas_cumprod = np.log(1. - alphas_cumprod)
        self.sqrt_recip_alphas_cumprod = np.sqrt(1. / alphas_cumprod)
        self.sqrt_recipm1_alphas_cumprod = np.sqrt(1. / alphas_cumprod - 1)
        ddim_sigmas, ddim_alphas, ddim_alphas_prev = make_ddim_sampling_parameters(
            alphacums=alphas_cumprod, ddim_timesteps=self.ddim_timesteps, eta=ddim_eta, verbose=verbose)
        self.ddim_sigmas = ddim_sigmas
        self.ddim_alphas = ddim_alphas
        self.ddim_alphas_prev = ddim_alphas_prev
        self.ddim_sqrt_one_minus_alphas = np.sqrt(1. - ddim_alphas)

    def sample(self, S, batch_size, shape, conditioning=None, callback=None, normals_sequence=None,
               img_callback=None, quantize_x0=False, eta=0., mask=None, x0=None, temperature=1., noise_dropout=0.,
               score_corrector=None, corrector_kwargs=None, verbose=True, x_T=None, log_every_t=100,
               unconditional_guidance_scale=1., unconditional_conditioning=None, **kwargs):
        _warn_batch(conditioning, batch_size)
        self.make_schedule(ddim_num_steps=S, ddim_eta=eta, verbose=False)
        C, H, W = shape
        size = (batch_size, C, H, W)
        if verbose:
            print('Data shape for %s sampling is %s' % ("DDIM" if self.ddim else "PLMS", (size,)))
        return self.plms_sampling(conditioning, size, callback=callback, img_callback=img_callback,
                                  quantize_denoised=quantize_x0, mask=mask, x0=x0, noise_dropout=noise_dropout,
                                  temperature=temperature, score_corrector=score_corrector,
                                  corrector_kwargs=corrector_kwargs, x_T=x_T, log_every_t=log_every_t,
                                  unconditional_guidance_scale=unconditional_guidance_scale,
                                  unconditional_conditioning=unconditional_conditioning, verbose=verbose)

    def _tables(self, time_range, model_type):
        """One (coefficient row, model-input time) per UNet evaluation, in launch order."""
        total = len(time_range)
        ac = self.alphas_cumprod
        rows, times = [], []

        def v_of(step):   # v -> eps at the evaluated timestep: eps = sqrt(abar_t) v + sqrt(1 - abar_t) x
            return (np.sqrt(ac[step]), np.sqrt(np.float32(1.) - ac[step])) if model_type == 1 else None

        for i, step in enumerate(time_range):
            index = total - i - 1
            step_next = time_range[min(i + 1, total - 1)]
            coef = (self.ddim_alphas[index], self.ddim_alphas_prev[index], self.ddim_sigmas[index],
                    self.ddim_sqrt_one_minus_alphas[index])
            if self.ddim:
                w = [((1., 0., 0., 0.), step)]
            elif i == 0:   # pseudo improved Euler, plms.py:231-235
                w = [((1., 0., 0., 0.), step), ((.5, .5, 0., 0.), step_next)]
            elif i == 1:   # Adams-Bashforth 2 / 3 / 4, plms.py:236-244
                w = [((3 / 2, -1 / 2, 0., 0.), step)]
            elif i == 2:
                w = [((23 / 12, -16 / 12, 5 / 12, 0.), step)]
            else:
                w = [((55 / 24, -59 / 24, 37 / 24, -9 / 24), step)]
            for wk, st in w:
                rows.append(ops.plms_coef_row(wk, coef, v_of(int(st)), use_noise=self.ddim))
                times.append(float(int(st)))
        return rows, times

    def plms_sampling(self, cond, shape, x_T=None, ddim_use_original_steps=False, callback=None, timesteps=None,
                      quantize_denoised=False, mask=None, x0=None, img_callback=None, log_every_t=100,
                      temperature=1., noise_dropout=0., score_corrector=None, corrector_kwargs=None,
                      unconditional_guidance_scale=1., unconditional_conditioning=None, verbose=True):
        _check_unsupported(mask, x0, score_corrector, quantize_denoised, noise_dropout)
        if ddim_use_original_steps or timesteps is not None:
            raise NotImplementedError("ddim_use_original_steps / explicit timesteps are not used by txt2img")
        device = self.model.unet.device
        unet = _GuidedUNet(self.model, cond, unconditional_conditioning, unconditional_guidance_scale)
        plan = unet.plan(shape)
        plan.img.copy_(torch.randn(shape, device=device, dtype=torch.float32) if x_T is None else
                       x_T.to(device=device, dtype=torch.float32))
        time_range = np.flip(self.ddim_timesteps)
        total_steps = self.ddim_timesteps.shape[0]
        plan.load_tables(*self._tables(time_range, plan.model_type))
        if mask is not None:
            mask = mask.to(device=device, dtype=torch.float32).contiguous()
            x0 = x0.to(device=device, dtype=torch.float32).contiguous()
        if verbose:
            print("Running %s Sampling with %d timesteps" % ("DDIM" if self.ddim else "PLMS", total_steps))
        first = plan.img.clone()
        intermediates = {'x_inter': [first], 'pred_x0': [first]}
        for i, step in enumerate(time_range):
            index = total_steps - i - 1
            if mask is not None:  # plms.py:153-156: img = q_sample(x0, ts, randn) * mask + (1 - mask) * img
                ops.sampler_mask_blend(plan.img, x0, torch.randn(shape, device=device, dtype=torch.float32), mask,
                                       self.model.sqrt_alphas_cumprod[step],
                                       self.model.sqrt_one_minus_alphas_cumprod[step])  # model.q_sample, ddpm.py:180-182
            if self.ddim:
                if float(self.ddim_sigmas[index]) != 0.0:  # noise_like(...) * temperature, plms.py:222
                    plan.noise.normal_().mul_(temperature)
                plan.run("ddim")
            else:
                plan.run("first" if i == 0 else "multi")
            if callback:
                callback(i)
            if img_callback:
                img_callback(plan.pred_x0, i)
            if index % log_every_t == 0 or index == total_steps - 1:
                intermediates['x_inter'].append(plan.img.clone())
                intermediates['pred_x0'].append(plan.pred_x0.clone())
        return plan.img.clone(), intermediates


class DDIMSampler(PLMSSampler):
    """DDIM (eta-parameterised) with the PLMSSampler API; see the module docstring."""
    ddim = True


# ----------------------------------------------------------------------------------------------------------
# DPM-Solver++ (2M)
# ----------------------------------------------------------------------------------------------------------
def interpolate_fn(x, xp, yp):
    """dpm_solver.py:1126-1171 for one channel: piecewise-linear through the key points, outermost segments
    extended.  The reference finds the segment by sorting concat(x, xp); a binary search gives the same segment
    (and the same value when x coincides with a key point)."""
    x = np.atleast_1d(np.asarray(x, dtype=np.float32))
    K = xp.shape[0]
    idx = np.searchsorted(xp, x, side="left")
    s = np.where(idx == 0, 0, np.where(idx == K, K - 2, idx - 1))
    return (yp[s] + (x - xp[s]) * (yp[s + 1] - yp[s]) / (xp[s + 1] - xp[s])).astype(np.float32)


class NoiseScheduleVP:
    """dpm_solver.py:20-168 for schedule='discrete' (the only one DPMSolverSampler builds, sampler.py:75)."""

    def __init__(self, schedule='discrete', betas=None, alphas_cumprod=None, continuous_beta_0=0.1,
                 continuous_beta_1=20.):
        if schedule != 'discrete':
            raise ValueError("Unsupported noise schedule {}. Only 'discrete' is built by the sampler".format(schedule))
        if betas is not None:
            log_alphas = 0.5 * np.cumsum(np.log(1 - np.asarray(betas, dtype=np.float32)))
        else:
            assert alphas_cumprod is not None
            log_alphas = 0.5 * np.log(np.asarray(alphas_cumprod, dtype=np.float32))
        self.schedule = schedule
        self.total_N = len(log_alphas)
        self.T = 1.
        t = np.linspace(np.float32(0.), np.float32(1.), self.total_N + 1, dtype=np.float32)[1:]
        self.t_array = t.astype(np.float16).astype(np.float32)  # fp16 cast of dpm_solver.py:100-103
        self.log_alpha_array = log_alphas.astype(np.float32)

    def marginal_log_mean_coeff(self, t):
        return interpolate_fn(t, self.t_array, self.log_alpha_array)

    def marginal_alpha(self, t):
        return np.exp(self.marginal_log_mean_coeff(t))

    def marginal_std(self, t):
        return np.sqrt(1. - np.exp(2. * self.marginal_log_mean_coeff(t)))

    def marginal_lambda(self, t):
        log_mean_coeff = self.marginal_log_mean_coeff(t)
        log_std = 0.5 * np.log(1. - np.exp(2. * log_mean_coeff))
        return log_mean_coeff - log_std


class DPMSolverSampler(object):
    def __init__(self, model, **kwargs):
        self.model = model
        self.register_buffer('alphas_cumprod', model.alphas_cumprod)

    def register_buffer(self, name, attr):
        setattr(self, name, attr)

    def sample(self, S, batch_size, shape, conditioning=None, callback=None, normals_sequence=None,
               img_callback=None, quantize_x0=False, eta=0., mask=None, x0=None, temperature=1., noise_dropout=0.,
               score_corrector=None, corrector_kwargs=None, verbose=True, x_T=None, log_every_t=100,
               unconditional_guidance_scale=1., unconditional_conditioning=None, **kwargs):
        _warn_batch(conditioning, batch_size)
        C, H, W = shape
        size = (batch_size, C, H, W)
        device = self.model.unet.device
        img = torch.randn(size, device=device, dtype=torch.float32) if x_T is None else \
            x_T.to(device=device, dtype=torch.float32)
        ns = NoiseScheduleVP('discrete', alphas_cumprod=self.alphas_cumprod)
        # sampler.py:77-85 hard-codes model_type="noise"; "v" (dpm_solver.py:281-284) is selected by the model's
        # parameterization so that the 768-v configuration runs through the same wrapper arithmetic.
        unet = _GuidedUNet(self.model, conditioning, unconditional_conditioning, unconditional_guidance_scale)
        plan = unet.plan(size)
        order, steps = 2, S
        if steps < order:
            raise AssertionError("steps >= order")  # dpm_solver.py:1064
        t_0, t_T = 1. / ns.total_N, ns.T
        # get_time_steps('time_uniform'): fp32 linspace cast to fp16 (dpm_solver.py:414-415)
        timesteps = np.linspace(np.float32(t_T), np.float32(t_0), steps + 1, dtype=np.float32).astype(np.float16)
        plan.img.copy_(img.to(torch.float16).to(torch.float32))  # ops.Cast()(img, ms.float16), sampler.py:89
        rows, times = [], []
        for i in range(steps):
            s16 = timesteps[i]
            # get_model_input_time on the fp16 tensor: (t - 1/N) * 1000 (dpm_solver.py:256-265)
            t_input = np.float32((s16 - np.float16(1. / ns.total_N)) * np.float16(1000.))
            s = np.float32(s16)
            t = np.float32(timesteps[i + 1])
            alpha_s, sigma_s = ns.marginal_alpha(s)[0], ns.marginal_std(s)[0]
            lam_s, lam_t = ns.marginal_lambda(s)[0], ns.marginal_lambda(t)[0]
            h = lam_t - lam_s
            sigma_t = ns.marginal_std(t)[0]
            alpha_t = np.exp(ns.marginal_log_mean_coeff(t)[0])
            step = i + 1  # index of the time we move to, as in the loop of dpm_solver.py:1073-1094
            if step == 1:
                step_order = 1
            elif steps < 15:  # lower_order_final
                step_order = min(order, steps + 1 - step)
            else:
                step_order = order
            if step_order == 1 or i == 0:
                # dpm_solver_first_update, predict_x0 branch (dpm_solver.py:504-513)
                coef = (alpha_s, sigma_s, sigma_t / sigma_s, -alpha_t * np.expm1(-h), 0., 0.)
            else:
                # multistep_dpm_solver_second_update, 'dpm_solver' type (dpm_solver.py:756-777)
                lam_p1 = ns.marginal_lambda(np.float32(timesteps[i - 1]))[0]
                h_0 = lam_s - lam_p1
                r0 = h_0 / h
                c = alpha_t * (np.exp(-h) - 1.)
                coef = (alpha_s, sigma_s, sigma_t / sigma_s, -c, -0.5 * c / r0, 1.)
            rows.append([float(np.float32(v)) for v in coef] + [0.] * (ops.STEP_COEFS - len(coef)))
            times.append(float(t_input))
        plan.load_tables(rows, times)
        for i in range(steps):
            plan.run("dpm")
            if callback:
                callback(i)
        return plan.img.clone(), None
