"""Times the d <= 64 tcgen05 attention of ONE library build (MDX_LIB picks it) on the UNet's shapes and reports its
error against an fp32 torch softmax(q k^T) v on the same bf16 inputs — used to A/B the share of exponentials moved
from the MUFU to the FMA pipe (builds with -DMDX_ATTN_POLY_NUM / _DEN).  Medians of 9 launches, CUDA events."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from minddiffusion_b200 import ops

tag = sys.argv[1] if len(sys.argv) > 1 else os.path.basename(os.environ.get("MDX_LIB", "default"))
SHAPES = [(128, 5, 4096, 4096, 64), (128, 10, 1024, 1024, 64), (128, 20, 256, 256, 64), (16, 8, 4096, 4096, 40), (128, 5, 4096, 77, 64)]
torch.manual_seed(0)
for B, heads, N, Nk, d in SHAPES:
    C = heads * d
    q = torch.randn(B, N, C, device="cuda").bfloat16()
    k = torch.randn(B, Nk, C, device="cuda").bfloat16()
    v = torch.randn(B, Nk, C, device="cuda").bfloat16()
    for _ in range(3):
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
    nb = 2
    qf, kf, vf = (t[:nb].float().view(nb, -1, heads, d).transpose(1, 2) for t in (q, k, v))
    ref = torch.softmax(qf @ kf.transpose(-1, -2) * d ** -0.5, -1) @ vf
    got = out[:nb].float().view(nb, N, heads, d).transpose(1, 2)
    err = ((got - ref).norm() / ref.norm()).item()
    print("%-14s B=%d h=%d Nq=%d Nk=%d d=%d %8.3f ms %6.0f TFLOP/s  rel-l2 err vs fp32 %.3e" % (
        tag, B, heads, N, Nk, d, ms[4], 4.0 * B * heads * N * Nk * d / ms[4] / 1e9, err), flush=True)
