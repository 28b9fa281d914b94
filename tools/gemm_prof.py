"""A few launches of one epilogue-bound GEMM shape for `ncu --set full --import-source on -k regex:gemm_tc`.
   python tools/gemm_prof.py [to_out|proj_in|geglu|qkv]"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from minddiffusion_b200 import ops

which = sys.argv[1] if len(sys.argv) > 1 else "to_out"
M, K = 524288, 320
N = {"to_out": 320, "proj_in": 320, "geglu": 2560, "qkv": 960}[which]
x = torch.randn(M, K, device="cuda").bfloat16()
w = (torch.randn(N, K, device="cuda") / K ** 0.5).bfloat16()
bias = torch.randn(N, device="cuda")
for _ in range(2):
    if which == "to_out":
        ops.gemm_tc([(x, 1)], w, bias=bias, residual=torch.randn(M, N, device="cuda"), out_dtype=torch.float32)
    elif which == "proj_in":
        ops.gemm_tc([(x, 1)], w, bias=bias, out_dtype=torch.float32)
    elif which == "geglu":
        ops.gemm_tc([(x, 1)], w, bias=bias, geglu=True)
    else:
        ops.gemm_tc([(x, 1)], w)
torch.cuda.synchronize()
print("ok")
