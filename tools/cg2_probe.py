"""Smallest check of the three launch forms of mdx_gemm_tc on one 3x3 convolution: plain, multicast pair, CTA pair
(tcgen05 cta_group::2) — each must run and be bit-identical to the first.  (This is the probe that showed a kernel
containing cta_group::2 instructions cannot be launched without a cluster: "cluster misconfiguration".)"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from minddiffusion_b200 import ops
n, H, W, Cin, Cout = 6, 64, 64, 64, 320
x = torch.randn(n, H, W, Cin, device="cuda").bfloat16()
w = (torch.randn(Cout, 9 * Cin, device="cuda") / (9 * Cin) ** 0.5).bfloat16()
ref = None
for name, env in [("plain", {"MDX_GEMM_MC": "0"}), ("mc2", {"MDX_GEMM_MC": "2"}), ("cg2", {"MDX_GEMM_MC": "2", "MDX_GEMM_CG2": "1"})]:
    os.environ.pop("MDX_GEMM_CG2", None)
    os.environ.update(env)
    try:
        out = ops.gemm_tc([(x, 9)], w, out_dtype=torch.float32)
        torch.cuda.synchronize()
        if ref is None:
            ref = out.clone()
        print(name, "ok", "equal" if torch.equal(out, ref) else "DIFF max %.3e" % float((out - ref).abs().max()), flush=True)
    except Exception as e:
        print(name, "FAILED", str(e)[:300], flush=True)
