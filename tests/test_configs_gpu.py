"""Parity at the BASELINE.json configuration shapes (-m gpu): the CUDA path, through the reference-shaped sampler API,
against full-size results of the CPU oracle committed as fixtures (tests/golden/fullsize_*.npz, written by
tests/golden/make_fullsize_oracle.py from the same seeded weights and inputs — the loops cost minutes of CPU time).

Tolerances are the north star's: per-evaluation max|diff| / max|ref| <= 1e-2 in bf16 and <= 1e-4 in the fp32 check
mode; final latents cosine >= 0.999.  They are asserted as stated, on single draws.
"""
import os

import numpy as np
import pytest
import torch

from conftest import ROOT, parity_log

pytestmark = pytest.mark.gpu

BF16_TOL, F32_TOL, COS = 1e-2, 1e-4, 0.999


def fixture(name):
    return dict(np.load(os.path.join(ROOT, "tests", "golden", "fullsize_%s.npz" % name)))


def rel_err(a, b):
    a = a.detach().float().cpu() if isinstance(a, torch.Tensor) else torch.from_numpy(np.asarray(a))
    b = b.detach().float().cpu() if isinstance(b, torch.Tensor) else torch.from_numpy(np.asarray(b))
    return float((a - b).abs().max() / (b.abs().max() + 1e-12))


def cosine(a, b):
    a = a.detach().double().cpu().flatten()
    b = torch.from_numpy(np.asarray(b)).double().flatten()
    return float(torch.dot(a, b) / (a.norm() * b.norm()))


def inputs(hw, ctx_dim, B=1):
    x_T = torch.from_numpy(np.random.RandomState(42).randn(B, 4, hw, hw).astype(np.float32))
    c = torch.from_numpy(np.random.RandomState(43).randn(B, 77, ctx_dim).astype(np.float32))
    uc = torch.from_numpy(np.random.RandomState(44).randn(B, 77, ctx_dim).astype(np.float32))
    return x_T.cuda(), c.cuda(), uc.cuda()


def make(cfg, dtype, **ldm_kw):
    from minddiffusion_b200.ldm import LatentDiffusion
    from minddiffusion_b200.unet import UNetModel
    from minddiffusion_b200.weights import random_unet_params
    model = UNetModel(random_unet_params(cfg, 0), dtype=dtype, **cfg)
    return model, LatentDiffusion(model, timesteps=1000, linear_start=0.00085, linear_end=0.012, **ldm_kw)


def raw_eval(model, x, t_value, uc, c):
    model.set_context(torch.cat((uc, c), 0))
    t = torch.full((2 * x.shape[0],), float(t_value), device="cuda")
    return model.forward_nhwc(x, t, repeat=2)[..., :4].permute(0, 3, 1, 2)


@pytest.mark.parametrize("dtype,tol", [(torch.float32, F32_TOL), (torch.bfloat16, BF16_TOL)], ids=["fp32", "bf16"])
def test_config3_sd2_64x64_ddim50_cfg(dtype, tol):
    """BASELINE config #3 / the bench shape: SD2 512-base UNet, 64x64 latent, 50-step DDIM, CFG 7.5.  One raw guided
    evaluation and the final latents of the whole 50-step loop, both against the CPU oracle's own full-size run."""
    from minddiffusion_b200 import samplers
    from oracle import sd_unet
    G = fixture("sd2_ddim50")
    model, ldm = make(sd_unet.SD2_UNET, dtype)
    x_T, c, uc = inputs(64, 1024)
    e = rel_err(raw_eval(model, x_T, 981.0, uc, c), G["raw_eps0"])
    parity_log("config3_sd2_64 eps %s" % dtype, e, tol)
    assert e <= tol, e
    model.use_cuda_graph = True
    out, _ = samplers.DDIMSampler(ldm).sample(50, 1, (4, 64, 64), c, x_T=x_T, verbose=False, eta=0.0,
                                              unconditional_guidance_scale=7.5, unconditional_conditioning=uc)
    assert torch.isfinite(out).all()
    parity_log("config3_sd2_64 ddim50 final 1-cos %s" % dtype, 1 - cosine(out, G["final"]), 1 - COS)
    assert cosine(out, G["final"]) >= COS, cosine(out, G["final"])
    if dtype == torch.float32:
        assert rel_err(out, G["final"]) < 2e-3


@pytest.mark.parametrize("dtype,tol", [(torch.float32, F32_TOL), (torch.bfloat16, BF16_TOL)], ids=["fp32", "bf16"])
def test_config4_sd2_96x96_v_prediction_dpm20_cfg(dtype, tol):
    """BASELINE config #4: SD2 768-v shape — 96x96 latent, v-prediction, 20-step DPM-Solver++(2M), CFG 7.5
    (dpm_solver.py:281-284 "v" branch after the guidance combine, :958-1123 multistep)."""
    from minddiffusion_b200 import samplers
    from oracle import sd_unet
    G = fixture("sd2v_dpm20")
    model, ldm = make(sd_unet.SD2_UNET, dtype, parameterization="v")
    x_T, c, uc = inputs(96, 1024)
    x16 = x_T.to(torch.float16).to(torch.float32)   # sampler.py:89
    e = rel_err(raw_eval(model, x16, float(G["t_in0"]), uc, c), G["raw_out0"])
    parity_log("config4_sd2v_96 out %s" % dtype, e, tol)
    assert e <= tol, e
    model.use_cuda_graph = True
    out, _ = samplers.DPMSolverSampler(ldm).sample(20, 1, (4, 96, 96), c, x_T=x_T, verbose=False,
                                                   unconditional_guidance_scale=7.5, unconditional_conditioning=uc)
    assert torch.isfinite(out).all()
    parity_log("config4_sd2v_96 dpm20 final 1-cos %s" % dtype, 1 - cosine(out, G["final"]), 1 - COS)
    assert cosine(out, G["final"]) >= COS, cosine(out, G["final"])
    if dtype == torch.float32:
        assert rel_err(out, G["final"]) < 2e-3


@pytest.mark.parametrize("dtype,tol", [(torch.float32, F32_TOL), (torch.bfloat16, BF16_TOL)], ids=["fp32", "bf16"])
def test_config2_wukong_64x64_plms_cfg(dtype, tol):
    """BASELINE config #2: the wukong-huahua UNet at full size (v1-inference-chinese.yaml:21-37 — 8 heads, so head dims
    40 / 80 / 160 inside the network, conv1x1 projections, 768-wide context), 64x64 latent, guided PLMS (4 steps =
    5 evaluations including the pseudo improved Euler one)."""
    from minddiffusion_b200 import samplers
    from oracle import sd_unet
    G = fixture("wukong_plms")
    model, ldm = make(sd_unet.WUKONG_UNET, dtype)
    x_T, c, uc = inputs(64, 768)
    e = rel_err(raw_eval(model, x_T, 751.0, uc, c), G["raw_eps0"])
    parity_log("config2_wukong_64 eps %s" % dtype, e, tol)
    assert e <= tol, e
    out, _ = samplers.PLMSSampler(ldm).sample(4, 1, (4, 64, 64), c, x_T=x_T, verbose=False,
                                              unconditional_guidance_scale=7.5, unconditional_conditioning=uc)
    parity_log("config2_wukong_64 plms4 final 1-cos %s" % dtype, 1 - cosine(out, G["final"]), 1 - COS)
    assert cosine(out, G["final"]) >= COS, cosine(out, G["final"])
    if dtype == torch.float32:
        assert rel_err(out, G["final"]) < 1e-3


@pytest.mark.parametrize("dtype,tol", [(torch.float32, F32_TOL), (torch.bfloat16, BF16_TOL)], ids=["fp32", "bf16"])
def test_config5_glide_upsampler_256(dtype, tol):
    """BASELINE config #5: the shipped GLIDE upsampler options at their real size, 256x256 (text2im_model.py:214-238),
    one evaluation."""
    from minddiffusion_b200 import glide
    G = fixture("glide_up256")
    opts = glide.model_and_diffusion_upsample(dtype=dtype)
    P = glide.random_glide_params(opts, 6, seed=2)
    T = lambda k: torch.from_numpy(G[k]).cuda()
    out = glide.create_upsample_model(P, **opts)(T("x"), torch.tensor([801.0]).cuda(), T("low_res"), T("tok"), T("mask"))
    assert tuple(out.shape) == G["out"].shape
    e = rel_err(out, G["out"])
    parity_log("config5_glide_up256 out %s" % dtype, e, tol)
    assert e <= tol, e
