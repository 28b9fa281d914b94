"""Single-GPU timing of the BASELINE.json configurations that are not the headline bench line:

  #1  stablediffusionv2 UNet, 256x256 (32x32 latent), batch 1, 2-step sampler with CFG (the reference's CPU case)
  #2  wukong-huahua 512x512 (64x64 latent), batch 8, 50-step PLMS (51 evaluations), CFG 7.5
  #4  stablediffusionv2 768-v (96x96 latent, v-prediction), 20-step DPM-Solver++(2M), per-GPU share of batch 16 on
      8 GPUs (= 2) and the whole batch 16 on one GPU
  #5  Taichu-GLIDE cascade: see tools/glide_bench.py

    python tools/config_bench.py [--which 1 2 4]

Synthetic conditioning / latents, seeded random-init weights, CUDA-event timing of the second call.  Every
configuration is timed twice: with the whole sampler step captured in CUDA graphs (device coefficient tables, the
product path) and with every kernel launched from the host, so the launch-bound configurations show what the capture
buys.  One JSON line per configuration; fractions are of MEASURED_PEAKS.json's burst bf16 peak.
"""
import argparse
import json
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from minddiffusion_b200 import _lib, samplers  # noqa: E402
from minddiffusion_b200.ldm import LatentDiffusion  # noqa: E402
from minddiffusion_b200.txt2img import SD2_UNET, WUKONG_UNET  # noqa: E402
from minddiffusion_b200.unet import UNetModel  # noqa: E402
from minddiffusion_b200.weights import random_unet_params  # noqa: E402

FLOP = {"wukong64": 803.3e9, "sd2_96": 2149.1e9, "sd2_32": 181.1e9}  # BASELINE.md §3, per sample per forward
try:
    PEAK = json.load(open(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "MEASURED_PEAKS.json")))["bf16_tflops"]
except (OSError, KeyError, ValueError):
    PEAK = 1590.0   # B200_PROFILING.md fallback


def timed(fn):
    fn()
    torch.cuda.synchronize()
    n0 = _lib.launch_count()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    out = fn()
    e1.record()
    torch.cuda.synchronize()
    assert torch.isfinite(out).all()
    return e0.elapsed_time(e1), _lib.launch_count() - n0


def run(name, cfg, sampler_cls, B, latent, S, evals, flop, parameterization="eps", ctx_dim=1024):
    unet = UNetModel(random_unet_params(cfg, 0), dtype=torch.bfloat16, **cfg)
    ldm = LatentDiffusion(unet, timesteps=1000, linear_start=0.00085, linear_end=0.012, scale_factor=0.18215,
                          parameterization=parameterization)
    sampler = sampler_cls(ldm)
    rs = np.random.RandomState(0)
    x = torch.from_numpy(rs.randn(B, 4, latent, latent).astype(np.float32)).cuda()
    c = torch.from_numpy(rs.randn(B, 77, ctx_dim).astype(np.float32)).cuda()
    uc = torch.from_numpy(rs.randn(B, 77, ctx_dim).astype(np.float32)).cuda()
    fn = lambda: sampler.sample(S, B, (4, latent, latent), c, x_T=x, verbose=False, unconditional_guidance_scale=7.5,
                                unconditional_conditioning=uc, eta=0.0)[0]
    unet.use_cuda_graph = False
    ms_eager, launches = timed(fn)
    unet.use_cuda_graph = True
    ms, host_launches = timed(fn)
    tf = B * 2 * evals * flop / (ms / 1e3) / 1e12
    print(json.dumps({"config": name, "batch": B, "latent": latent, "sampler_steps": S, "unet_evals": evals,
                      "ms_per_call": ms, "images_per_sec": B / (ms / 1e3), "unet_eval_ms": ms / evals,
                      "algorithmic_tflops": tf, "frac_of_burst_peak": tf / PEAK, "burst_peak_tflops": PEAK,
                      "ms_per_call_host_launched": ms_eager, "kernels_per_call": launches,
                      "host_launches_with_graphs": host_launches}), flush=True)
    del unet, ldm, sampler
    torch.cuda.empty_cache()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--which", type=int, nargs="*", default=[1, 2, 4])
    a = ap.parse_args()
    if 1 in a.which:
        run("#1 SD2 256^2 B1 2-step PLMS CFG7.5 (3 evaluations)", SD2_UNET, samplers.PLMSSampler, 1, 32, 2, 3, FLOP["sd2_32"])
        run("#1' SD2 256^2 B1 50-step DDIM CFG7.5", SD2_UNET, samplers.DDIMSampler, 1, 32, 50, 50, FLOP["sd2_32"])
    if 2 in a.which:
        run("#2 wukong-huahua 512^2 B8 50-step PLMS CFG7.5", WUKONG_UNET, samplers.PLMSSampler, 8, 64, 50, 51,
            FLOP["wukong64"], ctx_dim=768)
    if 4 in a.which:
        for B in (2, 16):
            run("#4 SD2 768-v B%d 20-step DPM-Solver++(2M) CFG7.5" % B, SD2_UNET, samplers.DPMSolverSampler, B, 96,
                20, 20, FLOP["sd2_96"], parameterization="v")


if __name__ == "__main__":
    main()
