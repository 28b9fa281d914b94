"""Generate golden vectors by EXECUTING the unmodified reference sources (under /root/reference) on the torch-backed
`mindspore` stand-in of ms_shim.py.  Run in the build container (the GPU box has no /root/reference):

    python tests/golden/make_golden.py --family sd2      # -> tests/golden/sd2_golden.npz
    python tests/golden/make_golden.py --family wukong   # -> tests/golden/wukong_golden.npz

What runs from the reference, unmodified:
  ldm/modules/diffusionmodules/openaimodel.py  UNetModel (constructor loops + construct), ResBlock, Up/Downsample
  ldm/modules/attention.py                      SpatialTransformer, BasicTransformerBlock, CrossAttention, GEGLU
  ldm/modules/diffusionmodules/util.py          timestep_embedding, make_beta_schedule, make_ddim_*
  ldm/models/diffusion/ddpm.py                  DDPM.register_schedule, LatentDiffusion.apply_model, DiffusionWrapper
  ldm/models/diffusion/plms.py                  PLMSSampler.sample (whole loop)
  ldm/models/diffusion/dpm_solver/*             DPMSolverSampler.sample, NoiseScheduleVP, model_wrapper, DPM_Solver
Weights are the seeded synthetic parameters of minddiffusion_b200.weights.random_unet_params (regenerated from the
seed by the tests, so only inputs' seeds and outputs are stored).  use_fp16=False: fp32 parameters and activations;
the fp16 roundings DPM-Solver hard-codes (time grid, t_array, cast of x_T) stay as the reference has them.
"""
import argparse
import os
import sys
import types

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, HERE)

import ms_shim  # noqa: E402

REF = {"sd2": "/root/reference/vision/stablediffusionv2", "wukong": "/root/reference/vision/wukong-huahua"}

TINY = {
    "sd2": dict(image_size=32, in_channels=4, out_channels=4, model_channels=64, attention_resolutions=[4, 2, 1],
                num_res_blocks=2, channel_mult=[1, 2, 4, 4], num_head_channels=64, use_spatial_transformer=True,
                use_linear_in_transformer=True, transformer_depth=1, context_dim=128, use_checkpoint=True,
                legacy=False, use_fp16=False),
    "wukong": dict(image_size=32, in_channels=4, out_channels=4, model_channels=64, attention_resolutions=[4, 2, 1],
                   num_res_blocks=2, channel_mult=[1, 2, 4, 4], num_heads=8, use_spatial_transformer=True,
                   transformer_depth=1, context_dim=192, use_checkpoint=True, legacy=False, use_fp16=False),
}


def oracle_cfg(cfg):
    """The subset of constructor arguments the oracle / product take (same keys as the YAML)."""
    keep = ("in_channels", "out_channels", "model_channels", "attention_resolutions", "num_res_blocks", "channel_mult",
            "num_head_channels", "num_heads", "use_spatial_transformer", "use_linear_in_transformer",
            "transformer_depth", "context_dim", "legacy")
    out = {k: cfg[k] for k in keep if k in cfg}
    out.setdefault("num_head_channels", -1)
    out.setdefault("num_heads", -1)
    out.setdefault("use_linear_in_transformer", False)
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--family", choices=["sd2", "wukong"], required=True)
    a = ap.parse_args()
    ms = ms_shim.install()
    sys.path.insert(0, REF[a.family])
    from ldm.models.diffusion import ddpm as ref_ddpm
    from ldm.models.diffusion.dpm_solver import DPMSolverSampler
    from ldm.models.diffusion.dpm_solver import dpm_solver as ref_dpm
    from ldm.models.diffusion.plms import PLMSSampler
    from ldm.modules.diffusionmodules import util as ref_util
    from minddiffusion_b200.weights import random_unet_params

    torch.manual_seed(0)
    torch.set_grad_enabled(False)
    cfg = TINY[a.family]
    ocfg = oracle_cfg(cfg)
    G = {}

    # ---- the model object the samplers consume, assembled from the reference's own pieces ----
    class Model:  # stands for LatentDiffusion without the VAE / text tower (not on the per-step path)
        pass

    model = Model()
    model.dtype = ms.float32
    model.v_posterior = 0.0
    model.parameterization = "eps"
    ref_ddpm.DDPM.register_schedule(model, given_betas=None, beta_schedule="linear", timesteps=1000,
                                    linear_start=0.00085, linear_end=0.012, cosine_s=8e-3)
    model.model = ref_ddpm.DiffusionWrapper(
        {"target": "ldm.modules.diffusionmodules.openaimodel.UNetModel", "params": cfg}, "crossattn")
    model.model.eval()
    apply_model = ref_ddpm.LatentDiffusion.apply_model
    record = []

    def recording_apply_model(self, *args, **kw):  # SD2 passes cond positionally, wukong as c_crossattn=
        out = apply_model(self, *args, **kw)
        record.append(out.detach().clone())
        return out

    model.apply_model = types.MethodType(recording_apply_model, model)

    # ---- seeded weights by MindSpore parameter name ----
    P = random_unet_params(ocfg, seed=0)
    named = dict(model.model.named_parameters())
    assert set("model." + k for k in named) == set(P), (set(P) ^ set("model." + k for k in named))
    for k, prm in named.items():
        v = torch.from_numpy(P["model." + k])
        assert tuple(v.shape) == tuple(prm.shape), (k, v.shape, prm.shape)
        prm.set_data(v)

    # ---- G1: schedules ----
    G["betas"] = model.betas.asnumpy()
    G["alphas_cumprod"] = model.alphas_cumprod.asnumpy()
    G["alphas_cumprod_prev"] = model.alphas_cumprod_prev.asnumpy()
    for S in (2, 50):
        G["ddim_timesteps_%d" % S] = ref_util.make_ddim_timesteps("uniform", S, 1000).asnumpy()
    sig, al, alp = ref_util.make_ddim_sampling_parameters(model.alphas_cumprod, ref_util.make_ddim_timesteps("uniform", 50, 1000), 0.0)
    G["ddim_alphas_50"], G["ddim_alphas_prev_50"], G["ddim_sigmas_50"] = al.asnumpy(), alp.asnumpy(), sig.asnumpy()
    ns = ref_dpm.NoiseScheduleVP("discrete", alphas_cumprod=model.alphas_cumprod)
    G["dpm_t_array"] = ns.t_array.asnumpy().astype(np.float32).reshape(-1)
    solver = ref_dpm.DPM_Solver(None, ns, predict_x0=True)
    for S in (6, 20):
        ts = solver.get_time_steps("time_uniform", 1.0, 1e-3, S, None)
        G["dpm_timesteps_%d" % S] = ts.asnumpy().astype(np.float32)
        G["dpm_lambda_%d" % S] = ns.marginal_lambda(ts).asnumpy().astype(np.float32)
        G["dpm_alpha_%d" % S] = ns.marginal_alpha(ts).asnumpy().astype(np.float32)
        G["dpm_std_%d" % S] = ns.marginal_std(ts).asnumpy().astype(np.float32)

    # ---- G2: one UNet evaluation ----
    rs = np.random.RandomState(1)
    x = rs.randn(2, 4, 16, 16).astype(np.float32)
    ctx = rs.randn(2, 77, cfg["context_dim"]).astype(np.float32)
    t = np.asarray([981, 21], dtype=np.int64)
    eps = model.model(ms.Tensor(x), ms.Tensor(t), c_crossattn=ms.Tensor(ctx))
    G["unet_eps"] = eps.asnumpy()
    G["unet_te"] = ref_util.timestep_embedding(ms.Tensor(t), cfg["model_channels"]).asnumpy()

    # ---- G3: PLMS, G4: DPM-Solver++ (both guided, scale 7.5) ----
    rs = np.random.RandomState(42)
    x_T = rs.randn(2, 4, 16, 16).astype(np.float32)
    c = rs.randn(2, 77, cfg["context_dim"]).astype(np.float32)
    uc = np.random.RandomState(44).randn(2, 77, cfg["context_dim"]).astype(np.float32)
    S = 6
    record.clear()
    out, inter = PLMSSampler(model).sample(S, 2, (4, 16, 16), ms.Tensor(c), verbose=False, x_T=ms.Tensor(x_T),
                                           unconditional_guidance_scale=7.5, unconditional_conditioning=ms.Tensor(uc))
    G["plms_out"] = out.asnumpy()
    G["plms_n_evals"] = np.asarray(len(record))
    G["plms_eval0"] = record[0].asnumpy()
    G["plms_eval1"] = record[1].asnumpy()
    G["plms_n_inter"] = np.asarray(len(inter["x_inter"]))
    # unguided (scale 1.0): the batch is not doubled
    record.clear()
    out, _ = PLMSSampler(model).sample(4, 2, (4, 16, 16), ms.Tensor(c), verbose=False, x_T=ms.Tensor(x_T))
    G["plms_unguided_out"] = out.asnumpy()
    G["plms_unguided_batch"] = np.asarray(record[0].shape[0])
    for S in (6, 20):
        record.clear()
        out, _ = DPMSolverSampler(model).sample(S, 2, (4, 16, 16), ms.Tensor(c), verbose=False, x_T=ms.Tensor(x_T),
                                                unconditional_guidance_scale=7.5,
                                                unconditional_conditioning=ms.Tensor(uc))
        G["dpm_out_%d" % S] = out.asnumpy().astype(np.float32)
        G["dpm_n_evals_%d" % S] = np.asarray(len(record))
    path = os.path.join(HERE, "%s_golden.npz" % a.family)
    np.savez_compressed(path, **G)
    print("wrote", path, {k: (v.shape, str(v.dtype)) for k, v in G.items() if v.ndim > 1})


if __name__ == "__main__":
    main()
