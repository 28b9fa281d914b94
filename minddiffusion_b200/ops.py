"""Operator layer: torch tensors in, sm_100a kernels (through the C ABI) out.

PyTorch is used for tensor ownership and the current CUDA stream only; every function here enqueues one or
more kernels of libmdx_b200.so.  Activations are NHWC: ``[n, H, W, C]`` or, flattened, ``[rows, C]``; the last
dim must be contiguous and the row stride uniform, so channel slices of wider buffers are valid operands.
"""
import ctypes

import torch

from . import _lib
from ._lib import STEP_COEFS, ConvArgs, GemmArgs, StepArgs, check

F32, BF16 = 0, 1
# bench.py sets this to a list to time every mdx_gemm_tc launch with CUDA events: entries (flops, start, end)
GEMM_TC_PROFILE = None
_DT = {torch.float32: F32, torch.bfloat16: BF16}


def _stream():
    return ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)


def _p(t):
    return ctypes.c_void_p(t.data_ptr()) if t is not None else None


def _dt(t):
    try:
        return _DT[t.dtype]
    except KeyError:
        raise TypeError("unsupported dtype %s (need float32 or bfloat16)" % t.dtype)


def _rows_ld(t):
    """(rows, C, ld) of an NHWC / [rows, C] view; raises unless rows are uniformly strided."""
    if not t.is_cuda:
        raise ValueError("expected a CUDA tensor (there is no CPU path)")
    if t.dim() < 2 or t.stride(-1) != 1:
        raise ValueError("last dim must be contiguous")
    ld = t.stride(-2)
    rows = t.size(-2)
    for d in range(t.dim() - 3, -1, -1):
        if t.size(d) != 1 and t.stride(d) != t.stride(d + 1) * t.size(d + 1):
            raise ValueError("rows of the view are not uniformly strided: %s %s" % (tuple(t.shape), t.stride()))
        rows *= t.size(d)
    return rows, t.size(-1), ld


def _f32(t, name):
    if t is None:
        return None
    if t.dtype != torch.float32 or not t.is_contiguous():
        raise TypeError("%s must be contiguous float32" % name)
    return t


def stat_slots(n_img, H, W, up=False):
    """Statistics slots per image of gemm_tc(chan_stats=...) for an [n_img, H, W, *] operand (0: unsupported)."""
    return int(_lib.load().mdx_gemm_tc_stat_slots(n_img, H, W, int(up)))


def pick_block_n(N, geglu=False):
    return int(_lib.load().mdx_gemm_tc_pick_block_n(N, 1 if geglu else 0))


ACT = {None: 0, "gelu": 1, "quick_gelu": 2, "tanh": 3}


def gemm_tc(segs, weight, out=None, bias=None, rowbias=None, residual=None, geglu=False, out_dtype=torch.bfloat16,
            block_n=0, act=None, up_phase=0, chan_stats=None, stride=1):
    """Tensor-core implicit GEMM (mdx_gemm_tc).

    segs: list of (a, taps); a is bf16 ``[n, H, W, C]`` (conv view) or ``[M, C]`` (dense), taps 9 or 1.
    weight: bf16 ``[N, Ktot]``.  rowbias: fp32 ``[n_img, N]`` view (row stride may exceed N).
    """
    lib = _lib.load()
    a0 = segs[0][0]
    if a0.dim() == 4:
        n_img, H, W = a0.shape[0], a0.shape[1], a0.shape[2]
    else:
        n_img, H, W = 1, 1, a0.shape[0]
    args = GemmArgs()
    args.nseg = len(segs)
    for i, (a, taps) in enumerate(segs):
        if a.dtype != torch.bfloat16:
            raise TypeError("gemm_tc operands must be bfloat16")
        rows, C, ld = _rows_ld(a)
        if rows != n_img * H * W:
            raise ValueError("segment %d has %d rows, expected %d" % (i, rows, n_img * H * W))
        args.a[i] = a.data_ptr()
        args.lda[i] = ld
        args.seg_c[i] = C
        args.seg_taps[i] = taps
    if weight.dtype != torch.bfloat16 or not weight.is_contiguous():
        raise TypeError("weight must be contiguous bfloat16 [N, K]")
    N = weight.shape[0]
    ktot = sum(taps * a.shape[-1] for a, taps in segs)
    if weight.shape[1] != ktot:
        raise ValueError("weight K=%d does not match segments K=%d" % (weight.shape[1], ktot))
    n_out = N // 2 if geglu else N
    OH, OW = ((H - 1) // 2 + 1, (W - 1) // 2 + 1) if stride == 2 else (H, W)   # stride 2: out is the strided conv's image
    M = n_img * OH * OW * (4 if up_phase else 1)   # up_phase: out / residual are the 2x-upsampled image
    if up_phase and out is None:
        raise ValueError("up_phase launches write one phase of a caller-provided [n, 2H, 2W, N] buffer")
    if out is None:
        out = torch.empty((M, n_out), dtype=out_dtype, device=a0.device)
    orow, oc, ldc = _rows_ld(out)
    if orow != M or oc != n_out:
        raise ValueError("out has shape %s, expected [%d, %d]" % (tuple(out.shape), M, n_out))
    args.n_img, args.H, args.W = n_img, H, W
    args.b = weight.data_ptr()
    args.N = N
    args.out = out.data_ptr()
    args.ldc = ldc
    args.out_dtype = _dt(out)
    args.bias = _f32(bias, "bias").data_ptr() if bias is not None else None
    if rowbias is not None:
        if rowbias.dtype != torch.float32 or rowbias.stride(-1) != 1:
            raise TypeError("rowbias must be float32 with contiguous last dim")
        args.rowbias = rowbias.data_ptr()
        args.ld_rowbias = rowbias.stride(0)
    if residual is not None:
        rr, rc, ldr = _rows_ld(residual)
        if rr != M or rc != n_out:
            raise ValueError("residual shape mismatch")
        args.residual = residual.data_ptr()
        args.ldr = ldr
        args.residual_dtype = _dt(residual)
    args.geglu = 1 if geglu else 0
    args.block_n = block_n
    args.act = ACT[act]
    args.up_phase = up_phase
    args.stride = stride
    if chan_stats is not None:
        slots = stat_slots(n_img, OH, OW, bool(up_phase))
        if slots == 0 or chan_stats.dtype != torch.float32 or not chan_stats.is_contiguous() or (
                tuple(chan_stats.shape) != (n_img, slots, 2, n_out)):
            raise ValueError("chan_stats must be contiguous float32 [n_img, %d, 2, %d]" % (slots, n_out))
        args.chan_stats = chan_stats.data_ptr()
    if GEMM_TC_PROFILE is not None:
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        check(lib.mdx_gemm_tc(ctypes.byref(args), _stream()), "mdx_gemm_tc")
        e1.record()
        GEMM_TC_PROFILE.append((2.0 * (M // (4 if up_phase else 1)) * N * ktot, e0, e1, (M, N, ktot, [t for _, t in segs], str(out.dtype)[6:],
                                                               None if residual is None else str(residual.dtype)[6:],
                                                               bool(geglu), rowbias is not None)))
        return out
    check(lib.mdx_gemm_tc(ctypes.byref(args), _stream()), "mdx_gemm_tc")
    return out


def conv_direct(x, w, kh=1, kw=1, stride=1, pad=0, upsample=False, pre_silu=False, bias=None, rowbias=None,
                residual=None, out=None, out_dtype=None, act=None):
    """CUDA-core implicit-GEMM conv/dense (mdx_conv_direct). x: [n,H,W,Cin] or [M,Cin]; w: [Cout, kh*kw*Cin]."""
    lib = _lib.load()
    if x.dim() == 4:
        n, H, W = x.shape[0], x.shape[1], x.shape[2]
    else:
        n, H, W = 1, 1, x.shape[0]
    rows, Cin, ldx = _rows_ld(x)
    Cout = w.shape[0]
    if w.dtype != x.dtype or not w.is_contiguous() or w.shape[1] != kh * kw * Cin:
        raise ValueError("weight must be contiguous [Cout, kh*kw*Cin] with the dtype of x")
    Hl, Wl = (2 * H, 2 * W) if upsample else (H, W)
    OH = (Hl + 2 * pad - kh) // stride + 1
    OW = (Wl + 2 * pad - kw) // stride + 1
    if out is None:
        odt = out_dtype or x.dtype
        out = torch.empty((n, OH, OW, Cout) if x.dim() == 4 else (OW, Cout), dtype=odt, device=x.device)
    orow, oc, ldo = _rows_ld(out)
    if orow != n * OH * OW or oc != Cout:
        raise ValueError("out shape mismatch")
    a = ConvArgs()
    a.x, a.ldx, a.in_dtype = x.data_ptr(), ldx, _dt(x)
    a.n, a.H, a.W, a.Cin = n, H, W, Cin
    a.w = w.data_ptr()
    a.Cout, a.kh, a.kw, a.stride, a.pad = Cout, kh, kw, stride, pad
    a.upsample, a.pre_silu = int(upsample), int(pre_silu)
    a.out, a.ldo, a.out_dtype = out.data_ptr(), ldo, _dt(out)
    a.bias = _f32(bias, "bias").data_ptr() if bias is not None else None
    if rowbias is not None:
        if rowbias.dtype != torch.float32 or rowbias.stride(-1) != 1:
            raise TypeError("rowbias must be float32 with contiguous last dim")
        a.rowbias, a.ld_rowbias = rowbias.data_ptr(), rowbias.stride(0)
    if residual is not None:
        if residual.dtype != out.dtype:
            raise TypeError("residual must have the out dtype")
        rr, rc, ldr = _rows_ld(residual)
        a.residual, a.ldr = residual.data_ptr(), ldr
    a.act = ACT[act]
    check(lib.mdx_conv_direct(ctypes.byref(a), _stream()), "mdx_conv_direct")
    return out


def groupnorm(x1, gamma, beta, groups=32, eps=1e-5, silu=False, x2=None, out=None, scratch=None, out_dtype=None,
              raw_copy=False, chan1=None, chan2=None):
    """GroupNorm(+SiLU) over NHWC ``[n, H, W, C]`` (optionally over concat(x1, x2) along C).
    out_dtype: dtype of the result (default: dtype of x1).  raw_copy=True also returns the un-normalised
    concat(x1, x2) converted to out_dtype (second return value)."""
    lib = _lib.load()
    n = x1.shape[0]
    rows, C1, ld1 = _rows_ld(x1)
    HW = rows // n
    C2, ld2 = 0, 0
    if x2 is not None:
        r2, C2, ld2 = _rows_ld(x2)
        if r2 != rows or x2.dtype != x1.dtype:
            raise ValueError("x2 must match x1 in rows and dtype")
    odt = out_dtype or x1.dtype
    if out is None:
        out = torch.empty(tuple(x1.shape[:-1]) + (C1 + C2,), dtype=odt, device=x1.device)
    _, oc, ldo = _rows_ld(out)
    if oc != C1 + C2:
        raise ValueError("out channel mismatch")
    raw = torch.empty_like(out) if raw_copy else None
    if scratch is None:
        scratch = torch.empty(n * groups * 2, dtype=torch.float64, device=x1.device)
    elif scratch.dtype != torch.float64 or scratch.numel() < n * groups * 2:
        raise ValueError("scratch must hold n*groups*2 float64")
    if chan1 is not None and (x2 is None or chan2 is not None):
        # statistics were accumulated by the producers' GEMM epilogues: no pass over the activation
        check(lib.mdx_groupnorm_pre(_p(x1), ld1, C1, _p(chan1), chan1.shape[1], _p(x2), ld2, C2, _p(chan2),
                                    chan2.shape[1] if chan2 is not None else 0, n, HW, groups, float(eps),
                                    _p(_f32(gamma, "gamma")), _p(_f32(beta, "beta")), None, 0, int(silu), _p(out), ldo,
                                    _p(raw), C1 + C2 if raw_copy else 0, _dt(x1), _dt(out), _p(scratch), _stream()),
              "mdx_groupnorm_pre")
        return (out, raw) if raw_copy else out
    check(lib.mdx_groupnorm(_p(x1), ld1, C1, _p(x2), ld2, C2, n, HW, groups, float(eps), _p(_f32(gamma, "gamma")),
                            _p(_f32(beta, "beta")), int(silu), _p(out), ldo, _p(raw), C1 + C2 if raw_copy else 0,
                            _dt(x1), _dt(out), _p(scratch), _stream()), "mdx_groupnorm")
    return (out, raw) if raw_copy else out


def layernorm(x, gamma, beta, eps=1e-5, out=None, out_dtype=None):
    lib = _lib.load()
    rows, C, ldx = _rows_ld(x)
    if out is None:
        out = torch.empty(x.shape, dtype=out_dtype or x.dtype, device=x.device)
    _, _, ldo = _rows_ld(out)
    check(lib.mdx_layernorm(_p(x), ldx, rows, C, float(eps), _p(_f32(gamma, "gamma")), _p(_f32(beta, "beta")),
                            _p(out), ldo, _dt(x), _dt(out), _stream()), "mdx_layernorm")
    return out


def softmax_rows(x, scale=1.0, out_dtype=None):
    """Row softmax of a [rows, cols] view: softmax(x * scale) along the last dim (fp32 math)."""
    lib = _lib.load()
    rows, cols, ldx = _rows_ld(x)
    out = torch.empty((rows, cols), dtype=out_dtype or x.dtype, device=x.device)
    check(lib.mdx_softmax_rows(_p(x), ldx, rows, cols, float(scale), _p(out), cols, _dt(x), _dt(out), _stream()),
          "mdx_softmax_rows")
    return out


def embedding(table, pos, ids, dtype=torch.float32):
    """out[i] = table[ids[i]] + pos[i % L]; table/pos fp32, ids int32 [batch, L] -> [batch*L, width]."""
    lib = _lib.load()
    table, pos = _f32(table, "table"), _f32(pos, "pos")
    if ids.dtype != torch.int32 or not ids.is_contiguous() or not ids.is_cuda:
        raise TypeError("ids must be a contiguous int32 CUDA tensor")
    n, L, width = ids.numel(), pos.shape[0], table.shape[1]
    out = torch.empty((n, width), dtype=dtype, device=table.device)
    check(lib.mdx_embedding(_p(table), table.shape[0], width, _p(pos), L, _p(ids), n, _p(out), _DT[dtype], _stream()),
          "mdx_embedding")
    return out


def attention(q, k, v, heads, scale, out=None, causal=False):
    """q: [B, Nq, heads*d] view, k/v: [B, Nk, heads*d] views (dense over batch). Returns [B, Nq, heads*d]."""
    lib = _lib.load()
    B, Nq, HDm = q.shape
    Nk = k.shape[1]
    d = HDm // heads
    for t, nm, nn in ((q, "q", Nq), (k, "k", Nk), (v, "v", Nk)):
        if t.stride(-1) != 1 or (B > 1 and t.stride(0) != t.stride(1) * nn):
            raise ValueError("%s must be dense over batch with contiguous channels" % nm)
    if out is None:
        out = torch.empty((B, Nq, HDm), dtype=q.dtype, device=q.device)
    check(lib.mdx_attention(_p(q), q.stride(1), _p(k), k.stride(1), _p(v), v.stride(1), _p(out), out.stride(1), B,
                            heads, Nq, Nk, d, float(scale), _dt(q), int(causal), _stream()), "mdx_attention")
    return out


def geglu(x, out=None):
    lib = _lib.load()
    rows, C2, ldx = _rows_ld(x)
    I = C2 // 2
    if out is None:
        out = torch.empty(tuple(x.shape[:-1]) + (I,), dtype=x.dtype, device=x.device)
    _, _, ldo = _rows_ld(out)
    check(lib.mdx_geglu(_p(x), ldx, rows, I, _p(out), ldo, _dt(x), _stream()), "mdx_geglu")
    return out


def timestep_embedding(t, dim, max_period=10000.0):
    lib = _lib.load()
    t = _f32(t, "t")
    out = torch.empty((t.shape[0], dim), dtype=torch.float32, device=t.device)
    check(lib.mdx_timestep_embedding(_p(t), t.shape[0], dim, float(max_period), _p(out), _stream()),
          "mdx_timestep_embedding")
    return out


def nchw_to_nhwc(x, dtype, repeat=1, cpad=None):
    lib = _lib.load()
    x = _f32(x, "x")
    n, C, H, W = x.shape
    cp = cpad or C
    out = torch.empty((n * repeat, H, W, cp), dtype=dtype, device=x.device)
    check(lib.mdx_nchw_to_nhwc_pad(_p(x), n, C, H, W, repeat, cp, _p(out), _DT[dtype], _stream()), "mdx_nchw_to_nhwc")
    return out


def nhwc_to_nchw(x, C=None):
    """x: [n, H, W, ld] -> fp32 [n, C, H, W] taking the first C channels."""
    lib = _lib.load()
    n, H, W, ld = x.shape
    if not x.is_contiguous():
        raise ValueError("x must be contiguous")
    C = C or ld
    out = torch.empty((n, C, H, W), dtype=torch.float32, device=x.device)
    check(lib.mdx_nhwc_to_nchw_ld(_p(x), ld, n, C, H, W, _dt(x), _p(out), _stream()), "mdx_nhwc_to_nchw")
    return out


def upsample_nearest2x(x, out_dtype=None):
    lib = _lib.load()
    n, H, W, C = x.shape
    if not x.is_contiguous():
        raise ValueError("x must be contiguous")
    out = torch.empty((n, 2 * H, 2 * W, C), dtype=out_dtype or x.dtype, device=x.device)
    check(lib.mdx_upsample_nearest2x(_p(x), n, H, W, C, _p(out), _dt(x), _dt(out), _stream()),
          "mdx_upsample_nearest2x")
    return out


def concat_channels(x1, x2):
    lib = _lib.load()
    r1, C1, ld1 = _rows_ld(x1)
    r2, C2, ld2 = _rows_ld(x2)
    if r1 != r2 or x1.dtype != x2.dtype:
        raise ValueError("concat operands mismatch")
    out = torch.empty(tuple(x1.shape[:-1]) + (C1 + C2,), dtype=x1.dtype, device=x1.device)
    check(lib.mdx_concat_channels(_p(x1), ld1, C1, _p(x2), ld2, C2, r1, _p(out), C1 + C2, _dt(x1), _stream()),
          "mdx_concat_channels")
    return out


def im2col3x3(x, stride, out_dtype=None):
    lib = _lib.load()
    n, H, W, C = x.shape
    if not x.is_contiguous():
        raise ValueError("x must be contiguous")
    OH, OW = (H + 2 - 3) // stride + 1, (W + 2 - 3) // stride + 1
    out = torch.empty((n, OH, OW, 9 * C), dtype=out_dtype or x.dtype, device=x.device)
    check(lib.mdx_im2col3x3(_p(x), n, H, W, C, stride, _p(out), _dt(x), _dt(out), _stream()), "mdx_im2col3x3")
    return out


def cast(x, dtype):
    lib = _lib.load()
    if not x.is_contiguous():
        raise ValueError("x must be contiguous")
    out = torch.empty(x.shape, dtype=dtype, device=x.device)
    check(lib.mdx_cast(_p(x), _dt(x), _p(out), _DT[dtype], x.numel(), _stream()), "mdx_cast")
    return out


def _host_floats(vals):
    arr = (ctypes.c_float * len(vals))(*[float(v) for v in vals])
    return arr


def sampler_plms_step(eps_model, x, w, coef, cfg, scale, C, old=(None, None, None), noise=None, x_prev=None,
                      pred_x0=None, e_t_out=None):
    """eps_model: NHWC UNet output [cfg?2B:B, h, w, ld]; x: fp32 NCHW [B, C, h, w]. See mdx.h for the formulas."""
    lib = _lib.load()
    B, Cx, h, wd = x.shape
    if Cx != C:
        raise ValueError("channel mismatch")
    x = _f32(x, "x")
    if x_prev is None:
        x_prev = torch.empty_like(x)
    ld = eps_model.shape[-1]
    if not eps_model.is_contiguous() or eps_model.shape[0] != (2 * B if cfg else B):
        raise ValueError("eps_model must be contiguous NHWC with %d images" % (2 * B if cfg else B))
    check(lib.mdx_sampler_plms_step(_p(eps_model), ld, _dt(eps_model), int(cfg), float(scale), _p(x),
                                    _p(_f32(old[0], "old1")), _p(_f32(old[1], "old2")), _p(_f32(old[2], "old3")),
                                    ctypes.cast(_host_floats(w), ctypes.c_void_p),
                                    ctypes.cast(_host_floats(coef), ctypes.c_void_p), _p(_f32(noise, "noise")),
                                    _p(x_prev), _p(pred_x0), _p(e_t_out), B, C, h, wd, _stream()),
          "mdx_sampler_plms_step")
    return x_prev


def sampler_dpmpp2m_step(eps_model, x, coef, cfg, scale, C, model_type=0, x0_prev=None, x_next=None, x0_cur=None):
    lib = _lib.load()
    B, Cx, h, wd = x.shape
    x = _f32(x, "x")
    if x_next is None:
        x_next = torch.empty_like(x)
    if x0_cur is None:
        x0_cur = torch.empty_like(x)
    ld = eps_model.shape[-1]
    if not eps_model.is_contiguous() or eps_model.shape[0] != (2 * B if cfg else B):
        raise ValueError("eps_model must be contiguous NHWC with %d images" % (2 * B if cfg else B))
    check(lib.mdx_sampler_dpmpp2m_step(_p(eps_model), ld, _dt(eps_model), int(cfg), float(scale), int(model_type),
                                       _p(x), _p(_f32(x0_prev, "x0_prev")),
                                       ctypes.cast(_host_floats(coef), ctypes.c_void_p), _p(x_next), _p(x0_cur), B, C,
                                       h, wd, _stream()), "mdx_sampler_dpmpp2m_step")
    return x_next, x0_cur


def plms_coef_row(w, coef, v=None, use_noise=False):
    """One table row of the PLMS / DDIM step kernel (mdx_sampler_plms_coef_row): w = AB weights (4), coef =
    (a_t, a_prev, sigma_t, sqrt_one_minus_a_t), v = (v_a, v_s) for a v-prediction model.  Returns a list of floats."""
    lib = _lib.load()
    row = (ctypes.c_float * STEP_COEFS)()
    check(lib.mdx_sampler_plms_coef_row(ctypes.cast(_host_floats(w), ctypes.c_void_p),
                                        ctypes.cast(_host_floats(coef), ctypes.c_void_p),
                                        ctypes.cast(_host_floats(v), ctypes.c_void_p) if v is not None else None,
                                        int(use_noise), ctypes.cast(row, ctypes.c_void_p)),
          "mdx_sampler_plms_coef_row")
    return list(row)


def sampler_step_ex(kind, model_out, x, x_out, C, cfg, scale, table=None, counter=None, row=None, model_type=0,
                    x_model=None, hist=(None, None, None), noise=None, pred_x0=None, e_t_out=None, shift_hist=False):
    """Struct form of the step kernels (mdx_sampler_plms_step_ex / mdx_sampler_dpmpp2m_step_ex).  kind: "plms" or
    "dpm".  Per-step scalars: `row` (host list of STEP_COEFS floats) or (`table` fp32 [rows, STEP_COEFS], `counter`
    int32 [1]) on the device."""
    lib = _lib.load()
    B, Cx, h, wd = x.shape
    if Cx != C:
        raise ValueError("channel mismatch")
    if not model_out.is_contiguous() or model_out.shape[0] != (2 * B if cfg else B):
        raise ValueError("model_out must be contiguous NHWC with %d images" % (2 * B if cfg else B))
    a = StepArgs()
    a.model_out, a.ld_out, a.model_dtype = model_out.data_ptr(), model_out.shape[-1], _dt(model_out)
    a.cfg, a.scale, a.model_type = int(cfg), float(scale), int(model_type)
    f = lambda t, nm: _f32(t, nm).data_ptr() if t is not None else None
    a.x, a.x_model = f(x, "x"), f(x_model, "x_model")
    for i in range(3):
        a.hist[i] = f(hist[i], "hist")
    a.noise, a.x_out, a.pred_x0, a.e_t_out = f(noise, "noise"), f(x_out, "x_out"), f(pred_x0, "pred_x0"), f(e_t_out, "e_t_out")
    a.shift_hist = int(shift_hist)
    a.B, a.C, a.h, a.w = B, C, h, wd
    keep = None
    if row is not None:
        if len(row) != STEP_COEFS:
            raise ValueError("row must hold %d floats" % STEP_COEFS)
        keep = _host_floats(row)
        a.coef_host = ctypes.cast(keep, ctypes.c_void_p)
    else:
        if table.dtype != torch.float32 or not table.is_contiguous() or table.shape[-1] != STEP_COEFS:
            raise ValueError("table must be contiguous float32 [rows, %d]" % STEP_COEFS)
        if counter.dtype != torch.int32 or not counter.is_cuda:
            raise ValueError("counter must be an int32 CUDA tensor")
        a.coef_table, a.step_dev = table.data_ptr(), counter.data_ptr()
    fn = lib.mdx_sampler_plms_step_ex if kind == "plms" else lib.mdx_sampler_dpmpp2m_step_ex
    check(fn(ctypes.byref(a), _stream()), "mdx_sampler_%s_step_ex" % kind)
    return x_out


def step_advance(counter, delta=1):
    check(_lib.load().mdx_step_advance(_p(counter), int(delta), _stream()), "mdx_step_advance")


def gather_row(table, counter, out, offset=0):
    """out[j, :] = table[counter + offset, :] for every row j of out (a 1-D out is n rows of width 1)."""
    t2 = table if table.dim() == 2 else table.view(-1, 1)
    o2 = out if out.dim() == 2 else out.view(-1, 1)
    if t2.dtype != torch.float32 or o2.dtype != torch.float32 or not o2.is_contiguous() or t2.stride(-1) != 1:
        raise TypeError("gather_row takes float32 tensors")
    check(_lib.load().mdx_gather_row(_p(t2), t2.stride(0), _p(counter), int(offset), _p(o2), o2.shape[0], o2.shape[1],
                                     _stream()), "mdx_gather_row")
    return out


def sampler_mask_blend(img, x0, noise, mask, sqrt_ac, sqrt_1m_ac):
    """img <- q_sample(x0, t, noise) * mask + (1 - mask) * img in place (plms.py:153-156); fp32 NCHW, mask [B, 1 or C, h, w]."""
    B, C, h, w = img.shape
    mask = _f32(mask, "mask")
    if mask.dim() != 4 or mask.shape[0] != B or mask.shape[2:] != img.shape[2:] or mask.shape[1] not in (1, C):
        raise ValueError("mask must be [B, 1 or C, h, w]")
    if x0.shape != img.shape or noise.shape != img.shape:
        raise ValueError("x0 / noise must have the latent's shape")
    check(_lib.load().mdx_sampler_mask_blend(_p(_f32(x0, "x0")), _p(_f32(noise, "noise")), _p(mask), mask.shape[1],
                                             float(sqrt_ac), float(sqrt_1m_ac), _p(_f32(img, "img")), B, C, h, w,
                                             _stream()), "mdx_sampler_mask_blend")
    return img


# ---------------------------------------------------------------------------------------------------------
# Taichu-GLIDE operators (include/mdx.h, "Taichu-GLIDE family")
# ---------------------------------------------------------------------------------------------------------
def groupnorm_film(x, gamma, beta, scale_shift, groups=32, eps=1e-5, silu=False, scratch=None, out_dtype=None):
    """GroupNorm(x) * (1 + scale) + shift (+SiLU); scale_shift: fp32 [n, 2C] view (row stride may exceed 2C)."""
    lib = _lib.load()
    n = x.shape[0]
    rows, C, ldx = _rows_ld(x)
    if scale_shift.dtype != torch.float32 or scale_shift.stride(-1) != 1 or scale_shift.shape != (n, 2 * C):
        raise ValueError("scale_shift must be float32 [n, 2C] with contiguous last dim")
    out = torch.empty(x.shape, dtype=out_dtype or x.dtype, device=x.device)
    if scratch is None:
        scratch = torch.empty(n * groups * 2, dtype=torch.float64, device=x.device)
    check(lib.mdx_groupnorm_film(_p(x), ldx, C, n, rows // n, groups, float(eps), _p(_f32(gamma, "gamma")),
                                 _p(_f32(beta, "beta")), _p(scale_shift), scale_shift.stride(0), int(silu), _p(out),
                                 C, _dt(x), _dt(out), _p(scratch), _stream()), "mdx_groupnorm_film")
    return out


def avgpool2x2(x, out_dtype=None):
    lib = _lib.load()
    n, H, W, C = x.shape
    if not x.is_contiguous():
        raise ValueError("x must be contiguous")
    out = torch.empty((n, H // 2, W // 2, C), dtype=out_dtype or x.dtype, device=x.device)
    check(lib.mdx_avgpool2x2(_p(x), n, H, W, C, _p(out), _dt(x), _dt(out), _stream()), "mdx_avgpool2x2")
    return out


def copy_rows(src, dst):
    """dst[b, r, :] = src[b, r, :] for 3-D views [nb, rows, width] with contiguous last dim; src.shape[0] == 1
    broadcasts."""
    lib = _lib.load()
    if src.dim() != 3 or dst.dim() != 3 or src.dtype != dst.dtype or src.stride(-1) != 1 or dst.stride(-1) != 1:
        raise ValueError("copy_rows takes 3-D views of one dtype with contiguous last dim")
    nb, rows, width = dst.shape
    if src.shape[1:] != dst.shape[1:] or src.shape[0] not in (1, nb):
        raise ValueError("copy_rows shape mismatch: %s -> %s" % (tuple(src.shape), tuple(dst.shape)))
    es = src.element_size()
    bs_src = 0 if (src.shape[0] == 1 and nb > 1) else src.stride(0) * es
    check(lib.mdx_copy_rows(_p(src), src.stride(1) * es, bs_src, _p(dst), dst.stride(1) * es, dst.stride(0) * es, nb,
                            rows, width * es, _stream()), "mdx_copy_rows")
    return dst


def embedding_masked(table, pos, pad, ids, mask=None, dtype=torch.float32):
    """out[i] = mask[i] ? table[ids[i]] + pos[i % L] : pad[i % L]; ids/mask int32 [batch, L]."""
    lib = _lib.load()
    table, pos = _f32(table, "table"), _f32(pos, "pos")
    for t, nm in ((ids, "ids"), (mask, "mask")):
        if t is not None and (t.dtype != torch.int32 or not t.is_contiguous() or not t.is_cuda):
            raise TypeError("%s must be a contiguous int32 CUDA tensor" % nm)
    n, L, width = ids.numel(), pos.shape[0], table.shape[1]
    out = torch.empty((n, width), dtype=dtype, device=table.device)
    check(lib.mdx_embedding_masked(_p(table), table.shape[0], width, _p(pos), _p(_f32(pad, "pad")), L, _p(ids),
                                   _p(mask), n, _p(out), _DT[dtype], _stream()), "mdx_embedding_masked")
    return out


def resize_bilinear(x, OH, OW, quantize_u8=False):
    lib = _lib.load()
    x = _f32(x, "x")
    n, C, H, W = x.shape
    out = torch.empty((n, C, OH, OW), dtype=torch.float32, device=x.device)
    check(lib.mdx_resize_bilinear(_p(x), n, C, H, W, OH, OW, int(quantize_u8), _p(out), _stream()),
          "mdx_resize_bilinear")
    return out


def nchw2_to_nhwc(x1, x2, dtype, repeat=1, cpad=None):
    """NHWC (channel-padded) image of concat(x1, x2) along channels; x2 may be None."""
    lib = _lib.load()
    x1 = _f32(x1, "x1")
    n, C1, H, W = x1.shape
    C2, n2 = 0, n
    if x2 is not None:
        x2 = _f32(x2, "x2")
        n2 = x2.shape[0]
        if n2 not in (n, n * repeat) or x2.shape[2:] != x1.shape[2:]:
            raise ValueError("x2 must hold n or n*repeat images of x1's size")
        C2 = x2.shape[1]
    cp = cpad or (C1 + C2)
    out = torch.empty((n * repeat, H, W, cp), dtype=dtype, device=x1.device)
    check(lib.mdx_nchw2_to_nhwc_pad(_p(x1), C1, _p(x2), C2, n, n2, H, W, repeat, cp, _p(out), _DT[dtype], _stream()),
          "mdx_nchw2_to_nhwc_pad")
    return out


def sampler_glide_step(model_out, x, coef, cfg, scale, mode, noise=None, x_prev=None, pred_xstart=None):
    """model_out: NHWC [cfg?2P:P, h, w, ld>=2C]; x: fp32 NCHW [P, C, h, w]; coef: 10 floats (mdx.h)."""
    lib = _lib.load()
    P, C, h, wd = x.shape
    x = _f32(x, "x")
    if x_prev is None:
        x_prev = torch.empty_like(x)
    if not model_out.is_contiguous() or model_out.shape[0] != (2 * P if cfg else P):
        raise ValueError("model_out must be contiguous NHWC with %d images" % (2 * P if cfg else P))
    check(lib.mdx_sampler_glide_step(_p(model_out), model_out.shape[-1], _dt(model_out), int(cfg), float(scale),
                                     int(mode), _p(x), ctypes.cast(_host_floats(coef), ctypes.c_void_p),
                                     _p(_f32(noise, "noise")), _p(x_prev), _p(pred_xstart), P, C, h, wd, _stream()),
          "mdx_sampler_glide_step")
    return x_prev


def prelu(x, slope=None, residual=None, out_dtype=None):
    """nn.PReLU(C) over NHWC rows (+ residual added after it); slope None = plain residual add."""
    lib = _lib.load()
    rows, C, ldx = _rows_ld(x)
    out = torch.empty(x.shape, dtype=out_dtype or x.dtype, device=x.device)
    ldr = 0
    if residual is not None:
        rr, rc, ldr = _rows_ld(residual)
        if rr != rows or rc != C or residual.dtype != x.dtype:
            raise ValueError("residual must match x")
    check(lib.mdx_prelu(_p(x), ldx, rows, C, _p(_f32(slope, "slope")), _p(residual), ldr, _p(out), C, _dt(x), _dt(out),
                        _stream()), "mdx_prelu")
    return out


def depth_to_space2(x, slope=None, out_dtype=None):
    """ops.DepthToSpace(2) on NHWC [n, H, W, 4C] -> [n, 2H, 2W, C], fused with a PReLU(C) when slope is given."""
    lib = _lib.load()
    n, H, W, C4 = x.shape
    if not x.is_contiguous() or C4 % 4:
        raise ValueError("x must be contiguous with a channel count divisible by 4")
    out = torch.empty((n, 2 * H, 2 * W, C4 // 4), dtype=out_dtype or x.dtype, device=x.device)
    check(lib.mdx_depth_to_space2(_p(x), n, H, W, C4 // 4, _p(_f32(slope, "slope")), _p(out), _dt(x), _dt(out), _stream()),
          "mdx_depth_to_space2")
    return out
