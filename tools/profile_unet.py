"""One guided SD2 UNet evaluation (CFG batch 2*B, 64x64 latent) for ncu: a warm-up forward, then one forward inside
the NVTX range "fwd" (ncu --nvtx --nvtx-include "fwd/").  Also prints CUDA-event time of that forward."""
import argparse
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch

from bench import SD2_UNET
from minddiffusion_b200.unet import UNetModel
from minddiffusion_b200.weights import random_unet_params

ap = argparse.ArgumentParser()
ap.add_argument("--batch", type=int, default=8)
ap.add_argument("--latent", type=int, default=64)
ap.add_argument("--iters", type=int, default=1)
a = ap.parse_args()
unet = UNetModel(random_unet_params(SD2_UNET, 0), dtype=torch.bfloat16, **SD2_UNET)
rs = np.random.RandomState(0)
x = torch.from_numpy(rs.randn(a.batch, 4, a.latent, a.latent).astype(np.float32)).cuda()
ctx = torch.from_numpy(rs.randn(2 * a.batch, 77, 1024).astype(np.float32)).cuda()
t = torch.full((2 * a.batch,), 501.0).cuda()
unet.set_context(ctx)
unet.forward_nhwc(x, t, repeat=2)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
torch.cuda.nvtx.range_push("fwd")
e0.record()
for _ in range(a.iters):
    out = unet.forward_nhwc(x, t, repeat=2)
e1.record()
torch.cuda.synchronize()
torch.cuda.nvtx.range_pop()
print("forward ms (eager, per iter): %.3f  finite=%s" % (e0.elapsed_time(e1) / a.iters, bool(torch.isfinite(out).all())))
