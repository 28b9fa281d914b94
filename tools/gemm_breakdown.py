"""Per-launch table of mdx_gemm_tc inside one guided SD2 UNet evaluation: shape, ms, TFLOP/s, algorithmic GB/s."""
import argparse
import os
import sys
from collections import OrderedDict

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch

from bench import SD2_UNET
from minddiffusion_b200 import ops
from minddiffusion_b200.unet import UNetModel
from minddiffusion_b200.weights import random_unet_params

ap = argparse.ArgumentParser()
ap.add_argument("--batch", type=int, default=64)
ap.add_argument("--latent", type=int, default=64)
a = ap.parse_args()
unet = UNetModel(random_unet_params(SD2_UNET, 0), dtype=torch.bfloat16, **SD2_UNET)
rs = np.random.RandomState(0)
x = torch.from_numpy(rs.randn(a.batch, 4, a.latent, a.latent).astype(np.float32)).cuda()
ctx = torch.from_numpy(rs.randn(2 * a.batch, 77, 1024).astype(np.float32)).cuda()
t = torch.full((2 * a.batch,), 501.0).cuda()
unet.set_context(ctx)
unet.forward_nhwc(x, t, repeat=2)
torch.cuda.synchronize()
prof = []
ops.GEMM_TC_PROFILE = prof
unet.forward_nhwc(x, t, repeat=2)
torch.cuda.synchronize()
ops.GEMM_TC_PROFILE = None
agg = OrderedDict()
for flops, e0, e1, (M, N, K, taps, odt, rdt, geglu, rb) in prof:
    ms = e0.elapsed_time(e1)
    ob = 4 if odt == "float32" else 2
    nout = N // 2 if geglu else N
    a_bytes = M * sum(K // sum(taps) * 1 for _ in [0]) * 2  # A read once (taps re-read from L2)
    a_bytes = M * (K // (9 if 9 in taps and len(taps) == 1 else 1)) * 2 if len(taps) == 1 else M * K * 2 // 9
    byts = a_bytes + N * K * 2 + M * nout * ob + (M * nout * (4 if rdt == "float32" else 2) if rdt else 0)
    key = (M, N, K, tuple(taps), odt, rdt, geglu)
    ent = agg.setdefault(key, [0, 0.0, flops, byts])
    ent[0] += 1
    ent[1] += ms
tot = sum(v[1] for v in agg.values())
print("%9s %6s %6s %-9s %-8s %-8s %5s | %4s %9s %7s %8s %8s" % ("M", "N", "K", "taps", "out", "resid", "geglu", "cnt", "ms(total)", "share", "TFLOP/s", "GB/s"))
for (M, N, K, taps, odt, rdt, geglu), (cnt, ms, flops, byts) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
    print("%9d %6d %6d %-9s %-8s %-8s %5s | %4d %9.3f %6.1f%% %8.0f %8.0f" % (
        M, N, K, ",".join(map(str, taps)), odt, rdt, geglu, cnt, ms, 100 * ms / tot, flops * cnt / ms / 1e9, byts * cnt / ms / 1e6))
print("total gemm_tc ms: %.3f over %d launches" % (tot, len(prof)))
