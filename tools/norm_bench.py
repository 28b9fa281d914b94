"""GroupNorm apply / LayerNorm at the SD2 UNet's shapes (CFG batch 128): achieved algorithmic GB/s, CUDA events,
L2 flushed before every timed launch.   python tools/norm_bench.py [label]"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from minddiffusion_b200 import ops

label = sys.argv[1] if len(sys.argv) > 1 else "cur"
dev = "cuda"
flush = torch.empty(512 << 20, dtype=torch.uint8, device=dev)


def timed(fn, reps=6):
    best = 1e9
    for _ in range(reps):
        flush.fill_(1)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        fn()
        e1.record()
        torch.cuda.synchronize()
        best = min(best, e0.elapsed_time(e1))
    return best


n = 128
for hw, C in ((64, 320), (32, 640), (16, 1280), (64, 640)):
    rows = n * hw * hw
    a32 = torch.randn(n, hw, hw, C, device=dev)
    a16 = a32.to(torch.bfloat16)
    g, b = torch.ones(C, device=dev), torch.zeros(C, device=dev)
    o16 = torch.empty(n, hw, hw, C, dtype=torch.bfloat16, device=dev)
    scratch = torch.empty(n * 64, dtype=torch.float64, device=dev)
    chan = torch.zeros((n, ops.stat_slots(n, hw, hw), 2, C), device=dev)
    chan[:, :, 1] = 1.0
    cases = [("gn f32->bf16 silu", rows * C * 6, lambda: ops.groupnorm(a32, g, b, 32, 1e-5, silu=True, out=o16, scratch=scratch, chan1=chan)),
             ("gn bf16->bf16 silu", rows * C * 4, lambda: ops.groupnorm(a16, g, b, 32, 1e-5, silu=True, out=o16, scratch=scratch, chan1=chan)),
             ("gn f32->bf16", rows * C * 6, lambda: ops.groupnorm(a32, g, b, 32, 1e-6, silu=False, out=o16, scratch=scratch, chan1=chan)),
             ("gn 2-pass f32->bf16 silu", rows * C * 10, lambda: ops.groupnorm(a32, g, b, 32, 1e-5, silu=True, out=o16, scratch=scratch)),
             ("ln f32->bf16", rows * C * 6, lambda: ops.layernorm(a32.view(rows, C), g, b, out=o16.view(rows, C)))]
    for name, nbytes, fn in cases:
        fn()
        ms = timed(fn)
        print("%-8s %dx%d C=%-5d %-26s %8.1f us %7.0f GB/s" % (label, hw, hw, C, name, ms * 1e3, nbytes / ms / 1e6), flush=True)
    del a32, a16, o16, chan
