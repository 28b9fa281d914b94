"""Sustained throughput of one kernel configuration under the power cap, A/B inside ONE process: alternate 4-second
bursts of back-to-back launches with an environment switch off / on and report launches per second of each burst.
   python tools/sustained_ab.py ENVVAR VALUE_ON [shape: conv|unet]"""
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from minddiffusion_b200 import ops

var, on = sys.argv[1], sys.argv[2]
n, hw, C = 128, 64, 320
x = torch.randn(n, hw, hw, C, device="cuda").bfloat16()
w = (torch.randn(C, 9 * C, device="cuda") / (9 * C) ** 0.5).bfloat16()
bias = torch.randn(C, device="cuda")
out = torch.empty(n * hw * hw, C, device="cuda", dtype=torch.bfloat16)
x2 = torch.randn(n, 32, 32, 640, device="cuda").bfloat16()
w2 = (torch.randn(640, 9 * 640, device="cuda") / (9 * 640) ** 0.5).bfloat16()
out2 = torch.empty(n * 32 * 32, 640, device="cuda", dtype=torch.bfloat16)


def burst(seconds):
    torch.cuda.synchronize()
    t0 = time.time()
    k = 0
    while time.time() - t0 < seconds:
        for _ in range(20):
            ops.gemm_tc([(x, 9)], w, out=out, bias=bias)
            ops.gemm_tc([(x2, 9)], w2, out=out2)
        torch.cuda.synchronize()
        k += 20
    return k / (time.time() - t0)


burst(3)
for rep in range(3):
    for val in (None, on):
        if val is None:
            os.environ.pop(var, None)
        else:
            os.environ[var] = val
        print("%s=%s: %.1f conv pairs/s" % (var, val, burst(4)), flush=True)
