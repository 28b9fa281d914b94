"""A/B of library variants on the small-K / residual GEMM shapes of the SD2 UNet (CFG batch 128): one line per
(variant, shape), CUDA events, L2 flushed between iterations.
   python tools/gemm_ab.py name=lib.so[:MDX_GEMM_DEBUG] ...      (lib relative to minddiffusion_b200/)"""
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
# (M, N, K per tap, taps, out, residual, geglu, label)
SHAPES = [(32768, 1280, 1280, 9, "bf16", None, False, "conv3x3 16^2 C1280"), (131072, 640, 640, 9, "bf16", None, False, "conv3x3 32^2 C640"), (524288, 320, 320, 1, "f32", "f32", False, "to_out 64^2"), (131072, 640, 640, 1, "f32", "f32", False, "to_out 32^2"),
          (32768, 1280, 1280, 1, "f32", "f32", False, "to_out 16^2"), (524288, 320, 1280, 1, "bf16", "f32", False, "ff2(last) 64^2"),
          (524288, 320, 320, 1, "f32", None, False, "proj_in 64^2"), (524288, 960, 320, 1, "bf16", None, False, "qkv 64^2"),
          (524288, 2560, 320, 1, "bf16", None, True, "geglu 64^2"), (131072, 5120, 640, 1, "bf16", None, True, "geglu 32^2"),
          (524288, 320, 320, 9, "f32", "f32", False, "conv3x3+res 64^2"), (524288, 320, 640, 9, "bf16", None, False, "conv3x3 K5760")]


def child():
    sys.path.insert(0, ROOT)
    import torch
    from minddiffusion_b200 import ops
    dt = {"f32": torch.float32, "bf16": torch.bfloat16}
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
    for M, N, K, taps, out_t, res_t, geglu, label in SHAPES:
        side = 16 if "16^2" in label else (32 if "32^2" in label else 64)
        x = (torch.randn(M // side ** 2, side, side, K, device="cuda") if taps == 9 else torch.randn(M, K, device="cuda")).bfloat16()
        w = (torch.randn(N, taps * K, device="cuda") / (taps * K) ** 0.5).bfloat16()
        bias = torch.randn(N, device="cuda")
        nout = N // 2 if geglu else N
        out = torch.empty(M, nout, device="cuda", dtype=dt[out_t])
        res = None if res_t is None else torch.randn(M, nout, device="cuda").to(dt[res_t])
        run = lambda: ops.gemm_tc([(x, taps)], w, out=out, bias=bias, residual=res, geglu=geglu)
        for _ in range(3):
            run()
        torch.cuda.synchronize()
        ms = []
        for _ in range(8):
            flush.zero_()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            run()
            e1.record()
            torch.cuda.synchronize()
            ms.append(e0.elapsed_time(e1))
        ms.sort()
        med = ms[len(ms) // 2]
        flops = 2.0 * M * N * taps * K
        byts = M * K * 2 + N * taps * K * 2 + M * nout * out.element_size() + (res.numel() * res.element_size() if res is not None else 0)
        print("%-12s %-18s %8.3f ms %6.0f TFLOP/s %6.0f GB/s" % (os.environ.get("AB_NAME", "?"), label, med, flops / med / 1e9, byts / med / 1e6), flush=True)
        del x, w, out, res


if __name__ == "__main__":
    if os.environ.get("AB_CHILD"):
        child()
    else:
        for v in sys.argv[1:] or ["cur=libmdx_b200.so"]:
            name, spec = v.split("=")
            lib, _, dbg = spec.partition(":")
            env = dict(os.environ, MDX_LIB=os.path.join(ROOT, "minddiffusion_b200", lib), AB_CHILD="1", AB_NAME=name)
            if dbg:
                env["MDX_GEMM_DEBUG"] = dbg
            r = subprocess.run([sys.executable, os.path.abspath(__file__)], env=env, capture_output=True, text=True)
            sys.stdout.write(r.stdout)
            if r.returncode:
                sys.stdout.write("%s FAILED: %s\n" % (name, r.stderr[-500:]))
            sys.stdout.flush()
