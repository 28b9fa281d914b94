"""Time mdx_attention (bf16) for one shape: python tools/attn_bench.py B heads Nq Nk d"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from minddiffusion_b200 import ops

B, heads, Nq, Nk, d = [int(v) for v in sys.argv[1:6]]
C = heads * d
qkv = torch.randn(B, Nq, 3 * C, device="cuda").bfloat16()
kv = torch.randn(B, Nk, 2 * C, device="cuda").bfloat16() if Nk != Nq else None
q = qkv[:, :, :C]
k = qkv[:, :, C:2 * C] if kv is None else kv[:, :, :C]
v = qkv[:, :, 2 * C:] if kv is None else kv[:, :, C:]
for _ in range(2):
    ops.attention(q, k, v, heads, d ** -0.5)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
iters = 5
for _ in range(iters):
    out = ops.attention(q, k, v, heads, d ** -0.5)
e1.record()
torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / iters
print("attention B=%d h=%d Nq=%d Nk=%d d=%d: %.3f ms  %.0f TFLOP/s  legacy=%s" % (
    B, heads, Nq, Nk, d, ms, 4.0 * B * heads * Nq * Nk * d / ms / 1e9, os.environ.get("MDX_ATTN_LEGACY")))
