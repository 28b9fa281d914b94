"""GroupNorm through the allocation-free entry point (-m gpu; collected last on purpose).

`mdx_groupnorm_noalloc` is for callers that reach the norm with no way to get device scratch — a framework
custom-op body that is given only tensor addresses and a stream, or a thread barred from the allocator — so the
per-group sums sit in a fixed array inside the library instead of a caller buffer.  The test drives it through
the C ABI with ctypes and checks it against the regular `mdx_groupnorm` entry point (same kernels, caller-owned
scratch) and against torch's own group_norm + SiLU on the concatenated input, twice in a row on different data: the
second call would pick up the first call's sums if the shared array were not cleared.  It also checks that a problem larger than the built-in array is refused with a clear
message rather than overrunning it.  Tolerances are the ones tests/test_ops_gpu.py uses for GroupNorm
(1e-4 fp32, 2e-2 bf16 of the output's max magnitude).
"""
import ctypes

import pytest
import torch

pytestmark = pytest.mark.gpu

MAX_PAIRS = 16384


def _call(lib, fn_name, x1, x2, gamma, beta, silu, eps, out, scratch=None):
    n = x1.shape[0]
    C1 = x1.shape[-1]
    C2 = x2.shape[-1] if x2 is not None else 0
    HW = x1.numel() // (n * C1)
    dt = {torch.float32: 0, torch.bfloat16: 1}
    p = lambda t: ctypes.c_void_p(t.data_ptr()) if t is not None else None
    stream = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
    args = [p(x1), C1, C1, p(x2), C2, C2, n, HW, 32, float(eps), p(gamma), p(beta), int(silu), p(out), C1 + C2,
            None, 0, dt[x1.dtype], dt[out.dtype]]
    if scratch is not None:
        args.append(p(scratch))
    args.append(stream)
    return getattr(lib, fn_name)(*args)


def _rel(a, b):
    a, b = a.float(), b.float()
    return float((a - b).abs().max() / (b.abs().max() + 1e-12))


@pytest.mark.parametrize("dtype,out_dtype,tol", [(torch.float32, torch.float32, 1e-4),
                                                 (torch.float32, torch.bfloat16, 2e-2),
                                                 (torch.bfloat16, torch.bfloat16, 2e-2)])
def test_noalloc_groupnorm_matches_regular_entry_point_twice(dtype, out_dtype, tol):
    from minddiffusion_b200 import _lib
    lib = _lib.load()
    g = torch.Generator().manual_seed(123)
    gamma = (1 + 0.1 * torch.randn(960, generator=g)).cuda()
    beta = (0.1 * torch.randn(960, generator=g)).cuda()
    for call, (shift, scale) in enumerate([(0.3, 1.5), (-2.0, 0.25)]):  # two calls in a row, different statistics
        x1 = (torch.randn(3, 10, 10, 640, generator=g) * scale + shift).cuda().to(dtype)
        x2 = (torch.randn(3, 10, 10, 320, generator=g) * scale - shift).cuda().to(dtype)
        want = torch.empty(3, 10, 10, 960, dtype=out_dtype, device="cuda")
        got = torch.full_like(want, float("nan"))
        scratch = torch.empty(3 * 32 * 2, dtype=torch.float64, device="cuda")
        assert _call(lib, "mdx_groupnorm", x1, x2, gamma, beta, True, 1e-5, want, scratch) == 0
        rc = _call(lib, "mdx_groupnorm_noalloc", x1, x2, gamma, beta, True, 1e-5, got)
        assert rc == 0, lib.mdx_last_error()
        torch.cuda.synchronize()
        assert torch.isfinite(got.float()).all()
        assert _rel(got, want) < tol, "call %d" % call
        # independent check: torch's GroupNorm + SiLU over concat(x1, x2) (what both entry points stand for)
        xin = torch.cat((x1, x2), dim=-1).float().permute(0, 3, 1, 2)
        ref = torch.nn.functional.silu(torch.nn.functional.group_norm(xin, 32, gamma, beta, 1e-5)).permute(0, 2, 3, 1)
        assert _rel(got, ref) < tol, "call %d vs torch" % call


def test_noalloc_groupnorm_rejects_sizes_beyond_the_builtin_scratch():
    from minddiffusion_b200 import _lib
    lib = _lib.load()
    n = MAX_PAIRS // 32 + 1  # one image too many for 32 groups
    x = torch.zeros(n, 1, 1, 64, device="cuda")
    out = torch.empty_like(x)
    gamma, beta = torch.ones(64, device="cuda"), torch.zeros(64, device="cuda")
    rc = _call(lib, "mdx_groupnorm_noalloc", x, None, gamma, beta, False, 1e-5, out)
    assert rc == -22
    assert b"exceeds the built-in scratch" in lib.mdx_last_error()
    # the largest size that fits is accepted
    x = torch.randn(MAX_PAIRS // 32, 1, 2, 64, device="cuda")
    out = torch.empty_like(x)
    assert _call(lib, "mdx_groupnorm_noalloc", x, None, gamma, beta, False, 1e-5, out) == 0
    torch.cuda.synchronize()
    ref = torch.nn.functional.group_norm(x.permute(0, 3, 1, 2), 32, gamma, beta, 1e-5).permute(0, 2, 3, 1)
    assert _rel(out, ref) < 1e-4
