"""Whole-evaluation A/B of an environment switch inside ONE process under the power cap: the guided SD2 UNet evaluation
(CFG batch 128, 64x64 latent) is run in alternating bursts with the switch unset / set; reports ms per evaluation of
each burst (CUDA events).  Launches are host-issued (no graph), which the GPU-bound evaluation hides.
   python tools/unet_env_ab.py ENVVAR VALUE [bursts] [evals per burst]"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch

from bench import SD2_UNET
from minddiffusion_b200.unet import UNetModel
from minddiffusion_b200.weights import random_unet_params

var, val = sys.argv[1], sys.argv[2]
bursts = int(sys.argv[3]) if len(sys.argv) > 3 else 4
per = int(sys.argv[4]) if len(sys.argv) > 4 else 12
os.environ["MDX_NO_FORWARD_GRAPH"] = "1"
unet = UNetModel(random_unet_params(SD2_UNET, 0), dtype=torch.bfloat16, **SD2_UNET)
rs = np.random.RandomState(0)
B = 64
x = torch.from_numpy(rs.randn(B, 4, 64, 64).astype(np.float32)).cuda()
ctx = torch.from_numpy(rs.randn(2 * B, 77, 1024).astype(np.float32)).cuda()
t = torch.full((2 * B,), 501.0).cuda()
unet.set_context(ctx)
for _ in range(6):
    unet._forward_impl(x, t, repeat=2) if hasattr(unet, "_forward_impl") else unet.forward_nhwc(x, t, repeat=2)
torch.cuda.synchronize()
res = {None: [], val: []}
for b in range(bursts):
    for v in (None, val):
        if v is None:
            os.environ.pop(var, None)
        else:
            os.environ[var] = v
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(per):
            unet._forward_impl(x, t, repeat=2)
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / per
        res[v].append(ms)
        print("%s=%s: %.3f ms per evaluation" % (var, v, ms), flush=True)
a, b = np.median(res[None]), np.median(res[val])
print("median: unset %.3f ms, %s=%s %.3f ms (%+.2f %%)" % (a, var, val, b, 100.0 * (b / a - 1.0)))
