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
