"""Kernel-level parity (-m gpu) for the operators the once-per-prompt networks add: causal attention (text tower),
GELU / QuickGELU GEMM epilogues, row softmax (VAE single-head attention), token embedding."""
import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu


def rel_err(a, b):
    a, b = a.detach().float().cpu(), b.detach().float().cpu()
    return float((a - b).abs().max() / (b.abs().max() + 1e-12))


def bf(x):
    return x.to(torch.bfloat16).float()


@pytest.mark.parametrize("dtype,tol", [(torch.float32, 1e-4), (torch.bfloat16, 2e-2)])
@pytest.mark.parametrize("B,heads,N,d", [(2, 16, 77, 64), (1, 12, 77, 64), (3, 4, 200, 32)])
def test_causal_attention(dtype, tol, B, heads, N, d):
    from minddiffusion_b200 import ops
    g = torch.Generator().manual_seed(N + d)
    C = heads * d
    qkv = torch.randn(B, N, 3 * C, generator=g)
    if dtype == torch.bfloat16:
        qkv = bf(qkv)
    q, k, v = qkv[..., :C], qkv[..., C:2 * C], qkv[..., 2 * C:]
    sp = lambda t: t.reshape(B, N, heads, d).permute(0, 2, 1, 3)
    s = torch.matmul(sp(q), sp(k).transpose(-1, -2)) * d ** -0.5
    s = s + torch.triu(torch.full((N, N), float("-inf")), 1)  # text_encoder.py:131-134
    ref = torch.matmul(torch.softmax(s, -1), sp(v)).permute(0, 2, 1, 3).reshape(B, N, C)
    dev = qkv.cuda().to(dtype)
    out = ops.attention(dev[..., :C], dev[..., C:2 * C], dev[..., 2 * C:], heads, d ** -0.5, causal=True)
    assert rel_err(out, ref) < tol


@pytest.mark.parametrize("act", ["gelu", "quick_gelu"])
def test_gemm_activation_epilogues(act):
    from minddiffusion_b200 import ops
    g = torch.Generator().manual_seed(5)
    M, K, N = 154, 256, 1024
    a = bf(torch.randn(M, K, generator=g))
    w = bf(torch.randn(N, K, generator=g) / K ** 0.5)
    b = torch.randn(N, generator=g)
    res = torch.randn(M, N, generator=g)
    y = a @ w.t() + b
    y = F.gelu(y, approximate="tanh") if act == "gelu" else y * torch.sigmoid(1.702 * y)
    out = ops.gemm_tc([(a.cuda().bfloat16(), 1)], w.cuda().bfloat16(), bias=b.cuda(), act=act, residual=res.cuda(),
                      out_dtype=torch.float32)
    assert rel_err(out, y + res) < 1e-3
    out32 = ops.conv_direct(a.cuda(), w.cuda(), bias=b.cuda(), act=act, residual=res.cuda())
    assert rel_err(out32, y + res) < 1e-4


@pytest.mark.parametrize("dtype,odt,tol", [(torch.float32, torch.float32, 1e-5), (torch.float32, torch.bfloat16, 1e-2),
                                           (torch.bfloat16, torch.bfloat16, 1e-2)])
@pytest.mark.parametrize("rows,cols", [(7, 4096), (130, 77), (3, 10)])
def test_softmax_rows(dtype, odt, tol, rows, cols):
    from minddiffusion_b200 import ops
    g = torch.Generator().manual_seed(rows + cols)
    x = (torch.randn(rows, cols, generator=g) * 4).to(dtype)
    ref = torch.softmax(x.float() * 0.37, -1)
    out = ops.softmax_rows(x.cuda(), 0.37, out_dtype=odt)
    assert out.dtype == odt and rel_err(out, ref) < tol


def test_embedding():
    from minddiffusion_b200 import ops
    g = torch.Generator().manual_seed(3)
    table, pos = torch.randn(1000, 64, generator=g), torch.randn(77, 64, generator=g)
    ids = torch.randint(0, 1000, (3, 77), generator=g, dtype=torch.int32)
    ref = (table[ids.long()] + pos[None]).reshape(-1, 64)
    out = ops.embedding(table.cuda(), pos.cuda(), ids.cuda())
    assert torch.equal(out.cpu(), ref)


@pytest.mark.gpu
@pytest.mark.parametrize("M,K,N,geglu", [(128 * 450 + 37, 320, 320, False), (128 * 230, 320, 960, False),
                                          (128 * 120, 192, 640, False), (128 * 60, 320, 2560, True),
                                          (128 * 500, 64, 128, False)])
def test_gemm_tc_resident_weights(M, K, N, geglu):
    """Small-K GEMMs keep one n-tile's weight slice in shared memory for the CTA's lifetime (gemm_tc.cu, resident
    mode).  Same inputs through the streaming schedule (MDX_GEMM_NO_RESIDENT=1) must give identical bits, and both
    must match torch."""
    import os
    from minddiffusion_b200 import ops
    from minddiffusion_b200.unet import pack_geglu
    g = torch.Generator().manual_seed(M % 1000 + K + N)
    a = torch.randn(M, K, generator=g).bfloat16()
    w = (torch.randn(N, K, generator=g) / K ** 0.5).bfloat16()
    bias = torch.randn(N, generator=g)
    res = torch.randn(M, N // 2 if geglu else N, generator=g)
    if geglu:
        val, gate = (a.float() @ w.float().t() + bias).chunk(2, dim=1)
        ref = val * torch.nn.functional.gelu(gate, approximate="tanh")
        wp, bp, bn = pack_geglu(w.cuda(), bias.cuda())
        run = lambda: ops.gemm_tc([(a.cuda(), 1)], wp, bias=bp, geglu=True, block_n=bn)
        tol = 2e-2
    else:
        ref = a.float() @ w.float().t() + bias + res
        run = lambda: ops.gemm_tc([(a.cuda(), 1)], w.cuda(), bias=bias.cuda(), residual=res.cuda(),
                                  out_dtype=torch.float32)
        tol = 1e-4
    out = run()
    os.environ["MDX_GEMM_NO_RESIDENT"] = "1"
    try:
        out_stream = run()
    finally:
        del os.environ["MDX_GEMM_NO_RESIDENT"]
    assert torch.equal(out, out_stream)
    err = float((out.float().cpu() - ref).abs().max() / ref.abs().max())
    assert err < tol, err


@pytest.mark.gpu
@pytest.mark.parametrize("n,H,W,Cin,Cout", [(2, 8, 8, 64, 64), (1, 16, 16, 128, 320), (3, 5, 7, 64, 128), (2, 32, 32, 64, 160)])
def test_upsample_conv_phase_folding(n, H, W, Cin, Cout):
    """Upsample (nearest x2) + conv3x3 as four 2x2 phase GEMMs over the source image (gemm_tc up_phase) against
    torch's interpolate + conv2d on the same bf16-rounded operands; also with a residual in the upsampled layout."""
    from minddiffusion_b200 import ops
    from minddiffusion_b200.unet import fold_upsample_weights, upsample_conv3x3
    g = torch.Generator().manual_seed(n * H + Cin + Cout)
    x = torch.randn(n, H, W, Cin, generator=g).bfloat16()
    w = (torch.randn(Cout, Cin, 3, 3, generator=g) / (9 * Cin) ** 0.5)
    b = torch.randn(Cout, generator=g)
    up = F.interpolate(x.float().permute(0, 3, 1, 2), scale_factor=2, mode="nearest")
    ref = F.conv2d(up, w, b, padding=1).permute(0, 2, 3, 1)
    wp = [p.cuda().bfloat16() for p in fold_upsample_weights(w)]
    out = upsample_conv3x3(x.cuda(), wp, b.cuda(), torch.float32)
    assert tuple(out.shape) == (n, 2 * H, 2 * W, Cout)
    err = float((out.cpu() - ref).abs().max() / ref.abs().max())
    assert err < 1e-2, err          # folded taps are rounded to bf16 once more than the 3x3 taps
    # exact fold: with weights that are exactly representable the phase GEMMs reproduce the 3x3 result to fp32 rounding
    wq = torch.randint(-2, 3, (Cout, Cin, 3, 3), generator=g).float() / 8
    ref = F.conv2d(up, wq, b, padding=1).permute(0, 2, 3, 1)
    res = torch.randn(n, 2 * H, 2 * W, Cout, generator=g)
    outq = torch.empty(n, 2 * H, 2 * W, Cout, device="cuda")
    for ph, p in enumerate(fold_upsample_weights(wq)):
        ops.gemm_tc([(x.cuda(), 4)], p.cuda().bfloat16(), out=outq.view(-1, Cout), bias=b.cuda(),
                    residual=res.cuda().view(-1, Cout), up_phase=ph + 1)
    err = float((outq.cpu() - (ref + res)).abs().max() / ref.abs().max())
    assert err < 1e-5, err


@pytest.mark.gpu
@pytest.mark.parametrize("B,heads,Nq,Nk", [(2, 5, 300, 77), (1, 3, 128, 128), (2, 2, 64, 1), (1, 4, 513, 16), (3, 2, 200, 100)])
def test_single_tile_attention_kernel(B, heads, Nq, Nk):
    """Nk <= 128 runs the small two-CTA-per-SM kernel (attention_tc.cu, attn_tc_single_kernel); it must agree with
    the general kernel (MDX_ATTN_NO_SINGLE=1) to bf16 rounding and with an fp32 torch softmax."""
    import os
    from minddiffusion_b200 import ops
    d, C = 64, heads * 64
    g = torch.Generator().manual_seed(B * Nq + Nk)
    q = torch.randn(B, Nq, C, generator=g).bfloat16()
    kv = torch.randn(B, Nk, 2 * C, generator=g).bfloat16()
    sp = lambda t, n: t.float().reshape(B, n, heads, d).permute(0, 2, 1, 3)
    s = torch.matmul(sp(q, Nq), sp(kv[..., :C], Nk).transpose(-1, -2)) * d ** -0.5
    ref = torch.matmul(torch.softmax(s, -1), sp(kv[..., C:], Nk)).permute(0, 2, 1, 3).reshape(B, Nq, C)
    kvd = kv.cuda()
    out = ops.attention(q.cuda(), kvd[..., :C], kvd[..., C:], heads, d ** -0.5)
    os.environ["MDX_ATTN_NO_SINGLE"] = "1"
    try:
        out_general = ops.attention(q.cuda(), kvd[..., :C], kvd[..., C:], heads, d ** -0.5)
    finally:
        del os.environ["MDX_ATTN_NO_SINGLE"]
    scale = float(ref.abs().max())
    assert float((out.float().cpu() - ref).abs().max()) / scale < 2e-2
    assert float((out.float() - out_general.float()).abs().max()) / scale < 1e-2


@pytest.mark.gpu
@pytest.mark.parametrize("n,H,W,Cin,C1,C2", [(3, 16, 16, 64, 320, 0), (2, 16, 8, 128, 640, 320), (2, 48, 40, 64, 64, 64),
                                              (2, 32, 32, 64, 160, 0)])
def test_groupnorm_from_epilogue_statistics(n, H, W, Cin, C1, C2):
    """The GEMM epilogue stores, per 32 output rows, the per-channel sums / sums of squares of what it wrote
    (chan_stats); mdx_groupnorm_pre folds them into group statistics.  Result must equal the two-pass GroupNorm of
    the same tensors (also for a concat whose groups straddle the two sources and for ragged tiles), and two runs
    must agree bit for bit (no atomics)."""
    from minddiffusion_b200 import ops
    g = torch.Generator().manual_seed(n + C1 + C2)
    slots = ops.stat_slots(n, H, W)
    assert slots > 0 and ops.stat_slots(4, 8, 8) == 0
    outs, chans = [], []
    for C in [c for c in (C1, C2) if c]:
        x = torch.randn(n, H, W, Cin, generator=g).bfloat16().cuda()
        w = (torch.randn(C, 9 * Cin, generator=g) / (9 * Cin) ** 0.5).bfloat16().cuda()
        res = torch.randn(n * H * W, C, generator=g).cuda()
        b = torch.randn(C, generator=g).cuda()
        chan = torch.full((n, slots, 2, C), float("nan"), device="cuda")
        o = ops.gemm_tc([(x, 9)], w, bias=b, residual=res, out_dtype=torch.float32, chan_stats=chan).view(n, H, W, C)
        ref = torch.stack((o.double().sum((1, 2)), (o.double() ** 2).sum((1, 2))), dim=1)
        assert torch.allclose(chan.double().sum(1), ref, rtol=1e-4, atol=1e-2)
        chan_b = torch.empty_like(chan)
        ops.gemm_tc([(x, 9)], w, bias=b, residual=res, out_dtype=torch.float32, chan_stats=chan_b)
        assert torch.equal(chan, chan_b)
        outs.append(o)
        chans.append(chan)
    C = C1 + C2
    gamma, beta = torch.randn(C, generator=g).cuda(), torch.randn(C, generator=g).cuda()
    x2, c2 = (outs[1], chans[1]) if C2 else (None, None)
    two_pass = ops.groupnorm(outs[0], gamma, beta, 32, 1e-5, silu=True, x2=x2, out_dtype=torch.bfloat16)
    fused = ops.groupnorm(outs[0], gamma, beta, 32, 1e-5, silu=True, x2=x2, out_dtype=torch.bfloat16, chan1=chans[0],
                          chan2=c2)
    assert float((two_pass.float() - fused.float()).abs().max()) <= 2e-2 * float(two_pass.float().abs().max())
    assert float((two_pass.float() - fused.float()).abs().mean()) < 1e-4
    # bf16 output: statistics are those of the rounded values
    chan = torch.empty(n, slots, 2, C1, device="cuda")
    x = torch.randn(n, H, W, Cin, generator=g).bfloat16().cuda()
    w = (torch.randn(C1, 9 * Cin, generator=g) / (9 * Cin) ** 0.5).bfloat16().cuda()
    o = ops.gemm_tc([(x, 9)], w, chan_stats=chan).view(n, H, W, C1)
    ref = torch.stack((o.double().sum((1, 2)), (o.double() ** 2).sum((1, 2))), dim=1)
    assert torch.allclose(chan.double().sum(1), ref, rtol=1e-4, atol=1e-2)
