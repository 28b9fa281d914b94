"""Full-size fixtures for the BASELINE.json configuration shapes, produced by the CPU ORACLE (oracle/*.py, itself
pinned against the executed reference by tests/test_oracle_golden.py) on the seeded weights the GPU tests rebuild.

    python tests/golden/make_fullsize_oracle.py [sd2_ddim50] [sd2v_dpm20] [wukong_plms] [glide_up256]

writes tests/golden/fullsize_<name>.npz.  The sampler loops take minutes of CPU time (100 full-size guided UNet
evaluations for the 50-step DDIM loop), which is why their results are committed instead of being recomputed inside
`pytest -m gpu` on the GPU box; the GPU tests (tests/test_configs_gpu.py) rebuild the same weights and inputs from
the same seeds and compare the CUDA path's per-evaluation eps and final latents with these files.

  sd2_ddim50   config #3 shape: SD2 512-base UNet, 64x64 latent, 1 image, 50-step DDIM, CFG 7.5
               (plms.py:210-228 update with e' = e_t), first evaluation's raw [uncond | cond] eps + final latents
  sd2v_dpm20   config #4 shape: SD2 UNet at a 96x96 latent, v-prediction, 20-step DPM-Solver++(2M), CFG 7.5
               (dpm_solver.py:281-284, 958-1123), first evaluation's raw output + final latents
  wukong_plms  config #2 shape: wukong-huahua UNet (8 heads: d = 40/80/160, conv1x1 projections, ctx 768), 64x64
               latent, 4-step PLMS, CFG 7.5 (5 evaluations incl. the pseudo improved Euler one), raw first eps + final
  glide_up256  config #5 shape: the shipped GLIDE upsampler options at 256x256, one evaluation
               (text2im_model.py:214-238)
"""
import os
import sys
import time

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)

from minddiffusion_b200.weights import random_unet_params  # noqa: E402  (pure numpy)
from oracle import samplers as osamp  # noqa: E402
from oracle import sd_unet  # noqa: E402


def inputs(hw, ctx_dim, B=1):
    """Seeds follow SURVEY.md §8d: x_T 42, cond 43, uncond 44."""
    x_T = torch.from_numpy(np.random.RandomState(42).randn(B, 4, hw, hw).astype(np.float32))
    c = torch.from_numpy(np.random.RandomState(43).randn(B, 77, ctx_dim).astype(np.float32))
    uc = torch.from_numpy(np.random.RandomState(44).randn(B, 77, ctx_dim).astype(np.float32))
    return x_T, c, uc


def unet_fn(cfg, seed=0):
    Pt = {k: torch.from_numpy(v) for k, v in random_unet_params(cfg, seed).items()}
    return Pt, (lambda x, t, c: sd_unet.unet_forward(Pt, cfg, x, t.to(torch.float32), c))


def sd2_ddim50():
    cfg = sd_unet.SD2_UNET
    Pt, fn = unet_fn(cfg)
    x_T, c, uc = inputs(64, 1024)
    sched = osamp.register_schedule(1000, 0.00085, 0.012)
    raw0 = sd_unet.unet_forward(Pt, cfg, torch.cat((x_T, x_T)), torch.full((2,), 981.0), torch.cat((uc, c)))
    rec = []
    out, _ = osamp.plms_sample(fn, sched, 50, x_T, c, uc, 7.5, ddim=True, record=rec)
    return dict(raw_eps0=raw0.numpy(), guided_eps0=rec[0].numpy(), guided_eps_last=rec[-1].numpy(), final=out.numpy())


def sd2v_dpm20():
    cfg = sd_unet.SD2_UNET
    Pt, fn = unet_fn(cfg)
    x_T, c, uc = inputs(96, 1024)
    sched = osamp.register_schedule(1000, 0.00085, 0.012)
    t_in = float(osamp.dpm_model_input_time(osamp.dpm_time_steps(20)[0]))
    x16 = x_T.to(torch.float16).to(torch.float32)
    raw0 = sd_unet.unet_forward(Pt, cfg, torch.cat((x16, x16)), torch.full((2,), t_in), torch.cat((uc, c)))
    rec = []
    out = osamp.dpmpp2m_sample(fn, sched, 20, x_T, c, uc, 7.5, model_type="v", record=rec)
    return dict(raw_out0=raw0.numpy(), t_in0=np.float32(t_in), guided_out0=rec[0].numpy(), final=out.numpy())


def wukong_plms():
    cfg = sd_unet.WUKONG_UNET
    Pt, fn = unet_fn(cfg)
    x_T, c, uc = inputs(64, 768)
    sched = osamp.register_schedule(1000, 0.00085, 0.012)
    raw0 = sd_unet.unet_forward(Pt, cfg, torch.cat((x_T, x_T)), torch.full((2,), 751.0), torch.cat((uc, c)))
    rec = []
    out, _ = osamp.plms_sample(fn, sched, 4, x_T, c, uc, 7.5, record=rec)
    assert len(rec) == 5
    return dict(raw_eps0=raw0.numpy(), guided_eps0=rec[0].numpy(), final=out.numpy())


def glide_up256():
    from minddiffusion_b200 import glide
    from oracle import glide as og
    opts = glide.model_and_diffusion_upsample()
    assert opts["image_size"] == 256
    P = glide.random_glide_params(opts, 6, seed=2)
    rs = np.random.RandomState(5)
    x = torch.from_numpy(rs.randn(1, 3, 256, 256).astype(np.float32))
    t = torch.tensor([801.0])
    tok = torch.from_numpy(rs.randint(1, opts["n_vocab"] - 1, size=(1, 128)).astype(np.int32))
    mask = torch.zeros(1, 128, dtype=torch.int32)
    mask[:, :21] = 1
    lr = torch.from_numpy(np.clip(rs.randn(1, 3, 64, 64) * 0.5, -1, 1).astype(np.float32))
    ref = og.unet_forward(P, opts, x, t, tok, mask, low_res=lr)
    return dict(x=x.numpy(), tok=tok.numpy(), mask=mask.numpy(), low_res=lr.numpy(), out=ref.numpy())


if __name__ == "__main__":
    torch.set_num_threads(os.cpu_count() or 1)
    todo = sys.argv[1:] or ["sd2_ddim50", "sd2v_dpm20", "wukong_plms", "glide_up256"]
    for name in todo:
        t0 = time.time()
        with torch.no_grad():
            res = globals()[name]()
        np.savez_compressed(os.path.join(HERE, "fullsize_%s.npz" % name), **res)
        print("%s: %.0f s, %s" % (name, time.time() - t0, {k: v.shape for k, v in res.items()}), flush=True)
