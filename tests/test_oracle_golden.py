"""CPU tier: the oracle against golden vectors produced by EXECUTING the unmodified reference sources
(tests/golden/make_golden.py runs /root/reference/vision/{stablediffusionv2,wukong-huahua}/ldm/... on the torch-backed
`mindspore` stand-in tests/golden/ms_shim.py; the reference has no tests or fixtures of its own).

Weights are regenerated from the seed (minddiffusion_b200.weights.random_unet_params), inputs from the RandomState
seeds recorded in make_golden.py.  Both sides are fp32 on CPU, so the tolerances are fp32 round-off: 1e-5 relative
for one UNet evaluation, 2e-4 for whole sampler trajectories (6-21 chained guided evaluations)."""
import os

import numpy as np
import pytest
import torch

from conftest import ROOT, TINY_UNET, TINY_UNET_V1

FAMILIES = {"sd2": TINY_UNET, "wukong": TINY_UNET_V1}


def golden(family):
    path = os.path.join(ROOT, "tests", "golden", "%s_golden.npz" % family)
    return dict(np.load(path))


def rel(a, b):
    a, b = np.asarray(a, dtype=np.float64), np.asarray(b, dtype=np.float64)
    return float(np.abs(a - b).max() / (np.abs(b).max() + 1e-30))


def params(cfg):
    from minddiffusion_b200.weights import random_unet_params
    return {k: torch.from_numpy(v) for k, v in random_unet_params(cfg, 0).items()}


@pytest.mark.parametrize("family", ["sd2", "wukong"])
def test_schedules_match_executed_reference(family):
    from oracle import samplers as osamp
    G = golden(family)
    s = osamp.register_schedule(1000, 0.00085, 0.012)
    assert np.allclose(s["betas"], G["betas"], rtol=1e-6, atol=0)  # LinSpace implementations differ by an ulp
    assert np.allclose(s["alphas_cumprod"], G["alphas_cumprod"], rtol=1e-6, atol=0)
    assert np.allclose(s["alphas_cumprod_prev"], G["alphas_cumprod_prev"], rtol=1e-6, atol=0)
    for S in (2, 50):
        assert np.array_equal(osamp.make_ddim_timesteps(S), G["ddim_timesteps_%d" % S])
    sig, al, alp = osamp.make_ddim_sampling_parameters(s["alphas_cumprod"], osamp.make_ddim_timesteps(50), 0.0)
    assert np.allclose(al, G["ddim_alphas_50"], rtol=1e-6) and np.allclose(alp, G["ddim_alphas_prev_50"], rtol=1e-6)
    assert np.array_equal(sig, G["ddim_sigmas_50"])
    ns = osamp.NoiseScheduleVP(G["alphas_cumprod"])
    assert np.array_equal(ns.t_array, G["dpm_t_array"])  # fp16-rounded grid, exactly
    for S in (6, 20):
        ts = osamp.dpm_time_steps(S).astype(np.float32)
        assert np.array_equal(ts, G["dpm_timesteps_%d" % S])
        assert np.allclose(ns.marginal_lambda(ts), G["dpm_lambda_%d" % S], rtol=2e-5, atol=2e-6)
        assert np.allclose(ns.marginal_alpha(ts), G["dpm_alpha_%d" % S], rtol=2e-5)
        assert np.allclose(ns.marginal_std(ts), G["dpm_std_%d" % S], rtol=2e-5)


@pytest.mark.parametrize("family", ["sd2", "wukong"])
def test_unet_matches_executed_reference(family):
    from oracle import sd_unet
    G, cfg = golden(family), FAMILIES[family]
    rs = np.random.RandomState(1)
    x = torch.from_numpy(rs.randn(2, 4, 16, 16).astype(np.float32))
    ctx = torch.from_numpy(rs.randn(2, 77, cfg["context_dim"]).astype(np.float32))
    t = torch.tensor([981, 21], dtype=torch.int64)
    assert rel(sd_unet.timestep_embedding(t, cfg["model_channels"]).numpy(), G["unet_te"]) < 1e-6
    eps = sd_unet.unet_forward(params(cfg), cfg, x, t, ctx)
    assert rel(eps.numpy(), G["unet_eps"]) < 1e-5


@pytest.mark.parametrize("family", ["sd2", "wukong"])
def test_samplers_match_executed_reference(family):
    from oracle import samplers as osamp
    from oracle import sd_unet
    G, cfg = golden(family), FAMILIES[family]
    P = params(cfg)
    rs = np.random.RandomState(42)
    x_T = torch.from_numpy(rs.randn(2, 4, 16, 16).astype(np.float32))
    c = torch.from_numpy(rs.randn(2, 77, cfg["context_dim"]).astype(np.float32))
    uc = torch.from_numpy(np.random.RandomState(44).randn(2, 77, cfg["context_dim"]).astype(np.float32))
    sched = osamp.register_schedule(1000, 0.00085, 0.012)
    calls = []

    def fn(x, t, cc):
        out = sd_unet.unet_forward(P, cfg, x, t.to(torch.float32), cc)
        calls.append(out)
        return out

    out, inter = osamp.plms_sample(fn, sched, 6, x_T, c, uc, 7.5)
    assert len(calls) == int(G["plms_n_evals"])  # S = 6 gives 7 timesteps (1000 // 6 = 166) + the extra first eval
    assert rel(calls[0].numpy(), G["plms_eval0"]) < 1e-5
    assert rel(calls[1].numpy(), G["plms_eval1"]) < 2e-5
    assert len(inter["x_inter"]) == int(G["plms_n_inter"])
    assert rel(out.numpy(), G["plms_out"]) < 2e-4
    calls.clear()
    out, _ = osamp.plms_sample(fn, sched, 4, x_T, c, None, 1.0)
    assert calls[0].shape[0] == int(G["plms_unguided_batch"]) == 2
    assert rel(out.numpy(), G["plms_unguided_out"]) < 2e-4
    for S in (6, 20):
        calls.clear()
        out = osamp.dpmpp2m_sample(fn, sched, S, x_T, c, uc, 7.5)
        assert len(calls) == int(G["dpm_n_evals_%d" % S]) == S
        assert rel(out.numpy(), G["dpm_out_%d" % S]) < 2e-4
