"""CPU tier (-m "not gpu"): the oracle's closed-form checks, host-side schedule logic against the oracle, and the
C-ABI library (loads, exports every symbol include/mdx.h declares; no compute without a GPU)."""
import ctypes
import os
import re

import numpy as np
import pytest
import torch

from conftest import ROOT, TINY_UNET, TINY_UNET_V1


def test_library_builds_loads_and_exports_every_declared_symbol():
    from minddiffusion_b200 import _lib
    lib = _lib.load()
    header = open(os.path.join(ROOT, "include", "mdx.h")).read()
    declared = set(re.findall(r"\b(mdx_[a-z0-9_]+)\s*\(", header))
    declared -= {"mdx_gemm_args", "mdx_conv_args", "mdx_step_args"}
    assert declared, "no declarations parsed"
    assert declared == set(_lib.SIGNATURES), declared ^ set(_lib.SIGNATURES)
    for name in declared:
        assert hasattr(lib, name), name
    assert lib.mdx_version() >= 100
    assert lib.mdx_gemm_tc_pick_block_n(320, 0) == 160
    assert lib.mdx_gemm_tc_pick_block_n(2560, 1) == 160


def test_ctypes_structs_match_the_compiled_header():
    """The argument structs of the binding (and of the stub in INTEGRATION.md) against the sizes the library was
    compiled with: a struct that is short by one field makes the library read that field from adjacent memory."""
    from minddiffusion_b200 import _lib
    lib = _lib.load()
    assert ctypes.sizeof(_lib.GemmArgs) == lib.mdx_sizeof_gemm_args()
    assert ctypes.sizeof(_lib.ConvArgs) == lib.mdx_sizeof_conv_args()
    assert ctypes.sizeof(_lib.StepArgs) == lib.mdx_sizeof_step_args()
    assert _lib.GemmArgs.chan_stats.offset == lib.mdx_offsetof_gemm_args_chan_stats()
    assert _lib.GemmArgs._fields_[-1][0] == "stride"
    # the documented stub declares the same fields in the same order
    doc = open(os.path.join(ROOT, "INTEGRATION.md")).read()
    stub = doc[doc.index("class GemmArgs"):doc.index("lib.mdx_gemm_tc.restype")]
    assert re.findall(r'\("([a-zA-Z_]+)", ctypes', stub) == [f[0] for f in _lib.GemmArgs._fields_]
    # plms coefficient rows: the host helper reproduces mdx_sampler_plms_step's arithmetic
    from minddiffusion_b200 import ops
    row = ops.plms_coef_row((1.5, -0.5, 0.0, 0.0), (0.25, 0.5, 0.0, 0.8660254), v=(0.5, 0.8660254))
    assert len(row) == _lib.STEP_COEFS and row[4] == 0.5 and abs(row[5] - 0.70710678) < 1e-7
    assert abs(row[6] - 0.70710678) < 1e-7 and row[9] == 0.5 and row[11] == 0.0


def test_abi_reports_errors_without_touching_the_gpu():
    from minddiffusion_b200 import _lib
    lib = _lib.load()
    rc = lib.mdx_gemm_tc(None, None)
    assert rc == -22 and b"NULL" in lib.mdx_last_error()
    rc = lib.mdx_layernorm(None, 0, 0, 0, 0.0, None, None, None, 0, 0, 0, None)
    assert rc == -22


@pytest.mark.skipif(torch.cuda.is_available(), reason="checks the no-GPU failure mode")
def test_product_path_fails_loudly_without_a_gpu():
    from minddiffusion_b200.unet import UNetModel
    from minddiffusion_b200.weights import random_unet_params
    with pytest.raises(RuntimeError, match="no CPU path"):
        UNetModel(random_unet_params(TINY_UNET, 0), **TINY_UNET)


def test_structure_and_param_inventory_match_oracle():
    from minddiffusion_b200.unet import unet_structure
    from minddiffusion_b200.weights import unet_param_shapes
    from oracle import sd_unet
    for cfg in (TINY_UNET, TINY_UNET_V1, sd_unet.SD2_UNET, sd_unet.WUKONG_UNET):
        assert unet_structure(cfg) == sd_unet.unet_structure(cfg)
        assert {k: tuple(v) for k, v in unet_param_shapes(cfg).items()} == \
               {k: tuple(v) for k, v in sd_unet.param_shapes(cfg).items()}
    inp, mid, out = unet_structure(sd_unet.SD2_UNET)
    layers = [l for b in inp + [mid] + out for l in b]
    # SURVEY.md §3.4 totals: 22 ResBlocks, 16 SpatialTransformers, 3 Down, 3 Up
    assert sum(l[0] == "res" for l in layers) == 22 and sum(l[0] == "st" for l in layers) == 16
    assert sum(l[0] == "down" for l in layers) == 3 and sum(l[0] == "up" for l in layers) == 3
    assert [l[2] for l in layers if l[0] == "st"][:3] == [5, 5, 10]  # heads = C / 64
    wk = [l for b in sd_unet.unet_structure(sd_unet.WUKONG_UNET)[0] for l in b if l[0] == "st"]
    assert [(l[2], l[3]) for l in wk] == [(8, 40), (8, 40), (8, 80), (8, 80), (8, 160), (8, 160)]
    n_params = sum(int(np.prod(s)) for s in unet_param_shapes(sd_unet.SD2_UNET).values())
    assert abs(n_params - 865.9e6) < 1e6


def test_schedule_known_answers():
    from oracle import samplers as osamp
    # make_ddim_timesteps arithmetic, util.py:135-145: S=50 -> 1,21,...,981 ; S=2 -> 1,501
    assert osamp.make_ddim_timesteps(50).tolist() == list(range(1, 1000, 20))
    assert osamp.make_ddim_timesteps(2).tolist() == [1, 501]
    s = osamp.register_schedule(1000, 0.00085, 0.012)
    b = s["betas"]
    assert b.shape == (1000,) and abs(b[0] - 0.00085) < 1e-9 and abs(b[-1] - 0.012) < 1e-8
    ref = np.cumprod(1 - np.linspace(0.00085 ** 0.5, 0.012 ** 0.5, 1000) ** 2)
    assert np.allclose(s["alphas_cumprod"], ref, rtol=2e-5)
    assert s["alphas_cumprod_prev"][0] == 1.0 and s["alphas_cumprod_prev"][1] == s["alphas_cumprod"][0]
    sig, a, ap = osamp.make_ddim_sampling_parameters(s["alphas_cumprod"], osamp.make_ddim_timesteps(50), 0.0)
    assert np.all(sig == 0) and ap[0] == s["alphas_cumprod"][0] and np.array_equal(ap[1:], a[:-1])


def test_host_schedules_match_oracle():
    from minddiffusion_b200 import ldm, samplers
    from oracle import samplers as osamp
    assert np.array_equal(ldm.make_beta_schedule("linear", 1000, 0.00085, 0.012), osamp.make_beta_schedule())
    for S in (2, 15, 20, 50):
        assert np.array_equal(samplers.make_ddim_timesteps("uniform", S, 1000), osamp.make_ddim_timesteps(S))
    ac = osamp.register_schedule()["alphas_cumprod"]
    for eta in (0.0, 0.7):
        got = samplers.make_ddim_sampling_parameters(ac, samplers.make_ddim_timesteps("uniform", 50), eta)
        want = osamp.make_ddim_sampling_parameters(ac, osamp.make_ddim_timesteps(50), eta)
        for g, w in zip(got, want):
            assert np.allclose(g, w, rtol=1e-6, atol=0)
    ns_a, ns_b = samplers.NoiseScheduleVP('discrete', alphas_cumprod=ac), osamp.NoiseScheduleVP(ac)
    ts = osamp.dpm_time_steps(20).astype(np.float32)
    assert np.allclose(ns_a.marginal_lambda(ts), ns_b.marginal_lambda(ts), rtol=1e-6)
    assert np.allclose(ns_a.marginal_std(ts), ns_b.marginal_std(ts), rtol=1e-6)
    with pytest.raises(ValueError):
        samplers.NoiseScheduleVP('linear')


def test_interpolate_fn_matches_sort_based_reference_form():
    """interpolate_fn restated literally with sort/argmin/gather (dpm_solver.py:1140-1170) vs the search form."""
    from oracle import samplers as osamp

    def literal(x, xp, yp):
        N, K = x.shape[0], xp.shape[0]
        all_x = np.concatenate((x[:, None], np.tile(xp[None], (N, 1))), axis=1)
        order = np.argsort(all_x, axis=1, kind="stable")
        sorted_all = np.take_along_axis(all_x, order, axis=1)
        x_idx = np.argmin(order, axis=1)
        cand = x_idx - 1
        start = np.where(x_idx == 0, 1, np.where(x_idx == K, K - 2, cand))
        end = np.where(start == cand, start + 2, start + 1)
        sx = np.take_along_axis(sorted_all, start[:, None], 1)[:, 0]
        ex = np.take_along_axis(sorted_all, end[:, None], 1)[:, 0]
        s2 = np.where(x_idx == 0, 0, np.where(x_idx == K, K - 2, cand))
        return yp[s2] + (x - sx) * (yp[s2 + 1] - yp[s2]) / (ex - sx)

    ac = osamp.register_schedule()["alphas_cumprod"]
    ns = osamp.NoiseScheduleVP(ac)
    rs = np.random.RandomState(0)
    x = np.concatenate((rs.uniform(-0.1, 1.1, 200), ns.t_array[[0, 5, 500, 999]], [0.001, 1.0])).astype(np.float32)
    a = osamp.interpolate_fn(x, ns.t_array, ns.log_alpha_array)
    b = literal(x, ns.t_array, ns.log_alpha_array)
    assert np.allclose(a, b, rtol=1e-5, atol=1e-7)


def test_oracle_samplers_closed_form():
    """With eps == 0 the DDIM/PLMS update is x * sqrt(a_prev / a_t) per step; with a linear eps model in x the
    DPM-Solver++ first update equals its closed form.  Guards the oracle against transcription slips."""
    from oracle import samplers as osamp
    sched = osamp.register_schedule()
    x = torch.randn(2, 4, 8, 8, generator=torch.Generator().manual_seed(0))
    zero = lambda x_, t, c: torch.zeros_like(x_)
    out, _ = osamp.plms_sample(zero, sched, 5, x, None)
    ts = osamp.make_ddim_timesteps(5)
    ac = sched["alphas_cumprod"].astype(np.float64)
    factor = np.sqrt(ac[0] / ac[ts[-1]])
    assert torch.allclose(out, x * float(factor), rtol=1e-4)
    out_d, _ = osamp.plms_sample(zero, sched, 5, x, None, ddim=True)
    assert torch.allclose(out_d, out, rtol=1e-5)
    # eps-prediction zero -> x0 = x / alpha ; one DPM++ step: x_t = (sigma_t/sigma_s) x - alpha_t expm1(-h) x/alpha_s
    ns = osamp.NoiseScheduleVP(sched["alphas_cumprod"])
    tsd = osamp.dpm_time_steps(2).astype(np.float32)
    got = osamp.dpmpp2m_sample(zero, sched, 2, x, None)
    s, t1, t2 = tsd
    def step1(xv, a, b):
        h = ns.marginal_lambda(b)[0] - ns.marginal_lambda(a)[0]
        return float(ns.marginal_std(b)[0] / ns.marginal_std(a)[0]) * xv - \
            float(ns.marginal_alpha(b)[0] * np.expm1(-h)) * (xv / float(ns.marginal_alpha(a)[0]))
    x16 = x.to(torch.float16).to(torch.float32)  # the reference rounds x_T to fp16 once (sampler.py:89)
    want = step1(step1(x16, s, t1), t1, t2)  # S=2 < 15: the last step falls back to order 1
    assert torch.allclose(got, want, rtol=1e-4, atol=1e-5)
    assert float(osamp.dpm_model_input_time(np.float16(1.0))) == 999.0  # fp16 rounding of (1 - 1/1000) * 1000


def test_oracle_unet_is_deterministic_and_uses_every_parameter():
    from minddiffusion_b200.weights import random_unet_params
    from oracle import sd_unet
    P = {k: torch.from_numpy(v) for k, v in random_unet_params(TINY_UNET_V1, 1).items()}
    x = torch.randn(1, 4, 8, 8, generator=torch.Generator().manual_seed(1))
    ctx = torch.randn(1, 7, TINY_UNET_V1["context_dim"], generator=torch.Generator().manual_seed(2))
    t = torch.tensor([10.0])
    y0 = sd_unet.unet_forward(P, TINY_UNET_V1, x, t, ctx)
    assert torch.equal(y0, sd_unet.unet_forward(P, TINY_UNET_V1, x, t, ctx))
    changed = 0
    names = sorted(P)[::7]
    for name in names:
        Q = dict(P)
        Q[name] = P[name] + 0.5
        changed += int(not torch.equal(sd_unet.unet_forward(Q, TINY_UNET_V1, x, t, ctx), y0))
    assert changed == len(names)


def test_lora_merge_equals_the_lora_forward():
    """W' = W + (alpha/rank) B A reproduces MindPet LoRADense's inference forward (weights.merge_lora)."""
    from minddiffusion_b200.weights import merge_lora
    rs = np.random.RandomState(0)
    stem = "model.diffusion_model.input_blocks.1.1.transformer_blocks.0.attn1.to_q."
    W, A, B = rs.randn(64, 48).astype(np.float32), rs.randn(4, 48).astype(np.float32), rs.randn(64, 4).astype(np.float32)
    x = rs.randn(5, 48).astype(np.float32)
    merged = merge_lora({stem + "weight": W, "other": W}, {stem + "mindpet_delta_lora_a": A, stem + "mindpet_delta_lora_b": B},
                        lora_rank=4, lora_alpha=8)
    want = x @ W.T + (8 / 4) * (x @ A.T) @ B.T
    np.testing.assert_allclose(x @ merged[stem + "weight"].T, want, rtol=1e-5, atol=1e-5)
    assert merged["other"] is W
    with pytest.raises(ValueError):
        merge_lora({stem + "weight": W}, {"unrelated": A})
    with pytest.raises(KeyError):
        merge_lora({stem + "weight": W}, {stem + "mindpet_delta_lora_a": A})


def test_mindspore_checkpoint_converter_round_trip():
    """tools/ckpt_to_npz.py: parser against a writer of the same checkpoint.proto schema (incl. sliced tensors,
    negative / large dims encoding, fp16 and bf16 payloads)."""
    import importlib.util
    spec = importlib.util.spec_from_file_location("ckpt_to_npz", os.path.join(ROOT, "tools", "ckpt_to_npz.py"))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    rs = np.random.RandomState(1)
    params = {"model.diffusion_model.time_embed.0.weight": rs.randn(7, 300).astype(np.float32),
              "first_stage_model.decoder.conv_in.bias": rs.randn(5).astype(np.float16),
              "step": np.array([12345678901], dtype=np.int64), "scalar": np.array(3.5, dtype=np.float32)}
    for slice_bytes in (None, 1000):
        got = mod.parse_checkpoint(mod.write_checkpoint(params, slice_bytes))
        assert list(got) == list(params)
        for k, v in params.items():
            assert got[k].dtype == v.dtype and got[k].shape == v.shape and np.array_equal(got[k], v)
    # bf16 payload: upper halves of fp32 words
    f = np.array([1.0, -2.5, 0.15625], dtype=np.float32)
    tensor = b"".join(mod._enc_varint(8) + mod._enc_varint(d) for d in (3,)) + mod._ld(2, b"BFloat16") + mod._ld(
        3, (f.view(np.uint32) >> 16).astype(np.uint16).tobytes())
    got = mod.parse_checkpoint(mod._ld(1, mod._ld(1, b"w") + mod._ld(2, tensor)))
    assert np.array_equal(got["w"], f)
