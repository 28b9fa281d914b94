"""A/B of the d <= 64 tcgen05 attention launch modes inside ONE process (same clocks, same buffers):
orig = attn_tc_kernel (one 128-query tile per CTA), static = one CTA per SM striding over all tiles, ipcN = N
consecutive tiles per CTA.  Times are medians of 9 launches, CUDA events."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from minddiffusion_b200 import ops

MODES = [("orig", {"MDX_ATTN_NO_PERSIST": "1"}), ("default", {}), ("static", {"MDX_ATTN_IPC": "0"}), ("ipc2", {"MDX_ATTN_IPC": "2"}),
         ("ipc3", {"MDX_ATTN_IPC": "3"}), ("ipc4", {"MDX_ATTN_IPC": "4"}), ("ipc6", {"MDX_ATTN_IPC": "6"})]
SHAPES = [(128, 5, 4096, 64), (128, 10, 1024, 64), (128, 20, 256, 64), (16, 8, 4096, 40), (32, 5, 9216, 64), (32, 5, 4096, 64)]
for B, heads, N, d in SHAPES:
    C = heads * d
    qkv = torch.randn(B, N, 3 * C, device="cuda").bfloat16()
    q, k, v = qkv[:, :, :C], qkv[:, :, C:2 * C], qkv[:, :, 2 * C:]
    ref = None
    for name, env in MODES:
        for key in ("MDX_ATTN_NO_PERSIST", "MDX_ATTN_IPC"):
            os.environ.pop(key, None)
        os.environ.update(env)
        for _ in range(2):
            out = ops.attention(q, k, v, heads, d ** -0.5)
        ms = []
        for _ in range(9):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            out = ops.attention(q, k, v, heads, d ** -0.5)
            e1.record()
            torch.cuda.synchronize()
            ms.append(e0.elapsed_time(e1))
        ms.sort()
        if ref is None:
            ref = out.clone()
        same = bool(torch.equal(out, ref))
        print("B=%d h=%d N=%d d=%d %-7s %8.3f ms %6.0f TFLOP/s  bit-equal-to-orig=%s" % (
            B, heads, N, d, name, ms[4], 4.0 * B * heads * N * N * d / ms[4] / 1e9, same), flush=True)
