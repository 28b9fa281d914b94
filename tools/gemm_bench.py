"""Time one mdx_gemm_tc shape in isolation (CUDA events, L2 flushed between iterations).
   python tools/gemm_bench.py M N K [--taps 9 --hw 64] [--out f32|bf16] [--res f32|bf16|none] [--geglu] [--bn N]"""
import argparse
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from minddiffusion_b200 import ops

ap = argparse.ArgumentParser()
ap.add_argument("M", type=int)
ap.add_argument("N", type=int)
ap.add_argument("K", type=int, help="channels per tap")
ap.add_argument("--taps", type=int, default=1)
ap.add_argument("--hw", type=int, default=64, help="image side when taps=9 (M = n*hw*hw)")
ap.add_argument("--out", default="bf16")
ap.add_argument("--res", default="none")
ap.add_argument("--geglu", action="store_true")
ap.add_argument("--bn", type=int, default=0)
ap.add_argument("--iters", type=int, default=10)
a = ap.parse_args()
dt = {"f32": torch.float32, "bf16": torch.bfloat16}
if a.taps == 9:
    n = a.M // (a.hw * a.hw)
    x = torch.randn(n, a.hw, a.hw, a.K, device="cuda").bfloat16()
else:
    x = torch.randn(a.M, a.K, device="cuda").bfloat16()
w = (torch.randn(a.N, a.taps * a.K, device="cuda") / (a.taps * a.K) ** 0.5).bfloat16()
bias = torch.randn(a.N, device="cuda")
nout = a.N // 2 if a.geglu else a.N
out = torch.empty(a.M, nout, device="cuda", dtype=dt[a.out])
res = None if a.res == "none" else torch.randn(a.M, nout, device="cuda").to(dt[a.res])
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
run = lambda: ops.gemm_tc([(x, a.taps)], w, out=out, bias=bias, residual=res, geglu=a.geglu, block_n=a.bn)
for _ in range(3):
    run()
torch.cuda.synchronize()
ms = []
for _ in range(a.iters):
    flush.zero_()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    run()
    e1.record()
    torch.cuda.synchronize()
    ms.append(e0.elapsed_time(e1))
ms.sort()
med = ms[len(ms) // 2]
flops = 2.0 * a.M * a.N * a.taps * a.K
byts = a.M * a.K * 2 + a.N * a.taps * a.K * 2 + a.M * nout * out.element_size() + (res.numel() * res.element_size() if res is not None else 0)
print("M=%d N=%d K=%dx%d out=%s res=%s geglu=%s bn=%d: %.3f ms  %.0f TFLOP/s  %.0f GB/s (algorithmic)" % (
    a.M, a.N, a.taps, a.K, a.out, a.res, a.geglu, a.bn, med, flops / med / 1e9, byts / med / 1e6))
