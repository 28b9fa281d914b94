"""In-process A/B of the cluster launch forms of mdx_gemm_tc (MDX_GEMM_MC is read per call): for every shape the
variants are timed alternately, several rounds, so that clock / power drift hits them equally.  CUDA events, L2 flushed
between iterations, medians."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from minddiffusion_b200 import ops

# (images, side, Cin, N, taps, out, residual, geglu, label)
SHAPES = [(128, 64, 320, 320, 9, "f32", "f32", False, "conv3x3+res 64^2 C320"), (128, 64, 640, 320, 9, "bf16", None, False, "conv3x3 64^2 K5760"),
          (128, 64, 960, 320, 9, "bf16", None, False, "conv3x3 64^2 K8640"), (128, 32, 640, 640, 9, "bf16", None, False, "conv3x3 32^2 C640"),
          (128, 32, 1920, 640, 9, "bf16", None, False, "conv3x3 32^2 K17280"), (128, 16, 1280, 1280, 9, "bf16", None, False, "conv3x3 16^2 C1280"),
          (128, 16, 2560, 1280, 9, "bf16", None, False, "conv3x3 16^2 K23040"), (128, 64, 320, 960, 1, "bf16", None, False, "qkv 64^2"),
          (128, 64, 320, 320, 1, "f32", "f32", False, "to_out 64^2"), (128, 64, 1280, 320, 1, "bf16", "f32", False, "ff2 64^2"),
          (128, 64, 320, 2560, 1, "bf16", None, True, "geglu 64^2"), (128, 32, 640, 5120, 1, "bf16", None, True, "geglu 32^2")]
VARIANTS = sys.argv[1:] or ["0", "2"]
dt = {"f32": torch.float32, "bf16": torch.bfloat16}
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
for n, side, K, N, taps, out_t, res_t, geglu, label in SHAPES:
    M = n * side * side
    x = (torch.randn(n, side, side, K, device="cuda") if taps == 9 else torch.randn(M, K, device="cuda")).bfloat16()
    w = (torch.randn(N, taps * K, device="cuda") / (taps * K) ** 0.5).bfloat16()
    bias = torch.randn(N, device="cuda")
    nout = N // 2 if geglu else N
    out = torch.empty(M, nout, device="cuda", dtype=dt[out_t])
    res = None if res_t is None else torch.randn(M, nout, device="cuda").to(dt[res_t])
    run = lambda: ops.gemm_tc([(x, taps)], w, out=out, bias=bias, residual=res, geglu=geglu)
    ms = {v: [] for v in VARIANTS}
    for rnd in range(4):
        for v in VARIANTS:
            if v in ("cg2", "halo"):            # CTA pairs (tcgen05 cta_group::2), without / with halo boxes
                os.environ["MDX_GEMM_MC"] = "2"
                os.environ["MDX_GEMM_CG2"] = "1"
                os.environ["MDX_GEMM_HALO"] = "1" if v == "halo" else "0"
            else:
                os.environ["MDX_GEMM_MC"] = v
                os.environ["MDX_GEMM_CG2"] = "0"
            run()
            for _ in range(5):
                flush.zero_()
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record()
                run()
                e1.record()
                torch.cuda.synchronize()
                ms[v].append(e0.elapsed_time(e1))
    flops = 2.0 * M * N * taps * K
    line = "%-24s" % label
    for v in VARIANTS:
        m = sorted(ms[v])[len(ms[v]) // 2]
        line += "  mc=%s %7.3f ms %5.0f TF" % (v, m, flops / m / 1e9)
    base = sorted(ms[VARIANTS[0]])[len(ms[VARIANTS[0]]) // 2]
    for v in VARIANTS[1:]:
        line += "  (%+.1f %%)" % (100.0 * (sorted(ms[v])[len(ms[v]) // 2] / base - 1.0))
    print(line, flush=True)
    del x, w, out, res
