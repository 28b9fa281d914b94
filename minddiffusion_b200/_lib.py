"""ctypes binding of libmdx_b200.so (include/mdx.h).

There is no CPU fallback: if the library cannot be built or loaded, importing this module raises, and every
wrapper raises `MdxError` when an entry point reports failure.
"""
import ctypes
import os

from . import build as _build

c_int = ctypes.c_int32
c_i64 = ctypes.c_int64
c_f = ctypes.c_float
c_p = ctypes.c_void_p


class MdxError(RuntimeError):
    pass


class GemmArgs(ctypes.Structure):
    _fields_ = [
        ("nseg", c_int),
        ("a", c_p * 3),
        ("lda", c_i64 * 3),
        ("seg_c", c_int * 3),
        ("seg_taps", c_int * 3),
        ("n_img", c_int), ("H", c_int), ("W", c_int),
        ("b", c_p),
        ("N", c_int),
        ("out", c_p),
        ("ldc", c_i64),
        ("out_dtype", c_int),
        ("bias", c_p),
        ("rowbias", c_p),
        ("ld_rowbias", c_i64),
        ("residual", c_p),
        ("ldr", c_i64),
        ("residual_dtype", c_int),
        ("geglu", c_int),
        ("block_n", c_int),
        ("act", c_int),
        ("up_phase", c_int),
        ("chan_stats", c_p),
        ("stride", c_int),
    ]


class ConvArgs(ctypes.Structure):
    _fields_ = [
        ("x", c_p), ("ldx", c_i64), ("in_dtype", c_int),
        ("n", c_int), ("H", c_int), ("W", c_int), ("Cin", c_int),
        ("w", c_p),
        ("Cout", c_int), ("kh", c_int), ("kw", c_int), ("stride", c_int), ("pad", c_int), ("upsample", c_int),
        ("pre_silu", c_int),
        ("out", c_p), ("ldo", c_i64), ("out_dtype", c_int),
        ("bias", c_p), ("rowbias", c_p), ("ld_rowbias", c_i64),
        ("residual", c_p), ("ldr", c_i64),
        ("act", c_int),
    ]


class StepArgs(ctypes.Structure):
    _fields_ = [
        ("model_out", c_p), ("ld_out", c_i64), ("model_dtype", c_int), ("cfg", c_int), ("scale", c_f),
        ("model_type", c_int),
        ("x", c_p), ("x_model", c_p),
        ("hist", c_p * 3),
        ("noise", c_p), ("x_out", c_p), ("pred_x0", c_p), ("e_t_out", c_p),
        ("shift_hist", c_int),
        ("B", c_int), ("C", c_int), ("h", c_int), ("w", c_int),
        ("coef_host", c_p), ("coef_table", c_p), ("step_dev", c_p),
    ]


STEP_COEFS = 16

# name -> (restype, argtypes); must list every symbol include/mdx.h declares (tests check this)
SIGNATURES = {
    "mdx_last_error": (ctypes.c_char_p, []),
    "mdx_version": (c_int, []),
    "mdx_launch_count": (c_i64, []),
    "mdx_gemm_tc": (c_int, [ctypes.POINTER(GemmArgs), c_p]),
    "mdx_gemm_tc_pick_block_n": (c_int, [c_int, c_int]),
    "mdx_conv_direct": (c_int, [ctypes.POINTER(ConvArgs), c_p]),
    "mdx_groupnorm": (c_int, [c_p, c_i64, c_int, c_p, c_i64, c_int, c_int, c_int, c_int, c_f, c_p, c_p, c_int, c_p,
                              c_i64, c_p, c_i64, c_int, c_int, c_p, c_p]),
    "mdx_groupnorm_noalloc": (c_int, [c_p, c_i64, c_int, c_p, c_i64, c_int, c_int, c_int, c_int, c_f, c_p, c_p, c_int,
                                      c_p, c_i64, c_p, c_i64, c_int, c_int, c_p]),
    "mdx_layernorm": (c_int, [c_p, c_i64, c_int, c_int, c_f, c_p, c_p, c_p, c_i64, c_int, c_int, c_p]),
    "mdx_attention": (c_int, [c_p, c_i64, c_p, c_i64, c_p, c_i64, c_p, c_i64, c_int, c_int, c_int, c_int, c_int, c_f,
                              c_int, c_int, c_p]),
    "mdx_softmax_rows": (c_int, [c_p, c_i64, c_i64, c_int, c_f, c_p, c_i64, c_int, c_int, c_p]),
    "mdx_embedding": (c_int, [c_p, c_int, c_int, c_p, c_int, c_p, c_int, c_p, c_int, c_p]),
    "mdx_geglu": (c_int, [c_p, c_i64, c_int, c_int, c_p, c_i64, c_int, c_p]),
    "mdx_timestep_embedding": (c_int, [c_p, c_int, c_int, c_f, c_p, c_p]),
    "mdx_nchw_to_nhwc": (c_int, [c_p, c_int, c_int, c_int, c_int, c_int, c_p, c_int, c_p]),
    "mdx_nchw_to_nhwc_pad": (c_int, [c_p, c_int, c_int, c_int, c_int, c_int, c_int, c_p, c_int, c_p]),
    "mdx_nhwc_to_nchw": (c_int, [c_p, c_int, c_int, c_int, c_int, c_int, c_p, c_p]),
    "mdx_nhwc_to_nchw_ld": (c_int, [c_p, c_i64, c_int, c_int, c_int, c_int, c_int, c_p, c_p]),
    "mdx_upsample_nearest2x": (c_int, [c_p, c_int, c_int, c_int, c_int, c_p, c_int, c_int, c_p]),
    "mdx_concat_channels": (c_int, [c_p, c_i64, c_int, c_p, c_i64, c_int, c_i64, c_p, c_i64, c_int, c_p]),
    "mdx_im2col3x3": (c_int, [c_p, c_int, c_int, c_int, c_int, c_int, c_p, c_int, c_int, c_p]),
    "mdx_cast": (c_int, [c_p, c_int, c_p, c_int, c_i64, c_p]),
    "mdx_sampler_plms_step": (c_int, [c_p, c_i64, c_int, c_int, c_f, c_p, c_p, c_p, c_p, c_p, c_p, c_p, c_p, c_p, c_p,
                                      c_int, c_int, c_int, c_int, c_p]),
    "mdx_sampler_dpmpp2m_step": (c_int, [c_p, c_i64, c_int, c_int, c_f, c_int, c_p, c_p, c_p, c_p, c_p, c_int, c_int,
                                         c_int, c_int, c_p]),
    "mdx_groupnorm_pre": (c_int, [c_p, c_i64, c_int, c_p, c_int, c_p, c_i64, c_int, c_p, c_int, c_int, c_int, c_int, c_f,
                                  c_p, c_p, c_p, c_i64, c_int, c_p, c_i64, c_p, c_i64, c_int, c_int, c_p, c_p]),
    "mdx_gemm_tc_stat_slots": (c_int, [c_int, c_int, c_int, c_int]),
    "mdx_groupnorm_film": (c_int, [c_p, c_i64, c_int, c_int, c_int, c_int, c_f, c_p, c_p, c_p, c_i64, c_int, c_p, c_i64,
                                   c_int, c_int, c_p, c_p]),
    "mdx_avgpool2x2": (c_int, [c_p, c_int, c_int, c_int, c_int, c_p, c_int, c_int, c_p]),
    "mdx_copy_rows": (c_int, [c_p, c_i64, c_i64, c_p, c_i64, c_i64, c_int, c_int, c_i64, c_p]),
    "mdx_embedding_masked": (c_int, [c_p, c_int, c_int, c_p, c_p, c_int, c_p, c_p, c_int, c_p, c_int, c_p]),
    "mdx_resize_bilinear": (c_int, [c_p, c_int, c_int, c_int, c_int, c_int, c_int, c_int, c_p, c_p]),
    "mdx_nchw2_to_nhwc_pad": (c_int, [c_p, c_int, c_p, c_int, c_int, c_int, c_int, c_int, c_int, c_int, c_p, c_int,
                                      c_p]),
    "mdx_prelu": (c_int, [c_p, c_i64, c_i64, c_int, c_p, c_p, c_i64, c_p, c_i64, c_int, c_int, c_p]),
    "mdx_depth_to_space2": (c_int, [c_p, c_int, c_int, c_int, c_int, c_p, c_p, c_int, c_int, c_p]),
    "mdx_sampler_plms_step_ex": (c_int, [ctypes.POINTER(StepArgs), c_p]),
    "mdx_sampler_dpmpp2m_step_ex": (c_int, [ctypes.POINTER(StepArgs), c_p]),
    "mdx_sampler_plms_coef_row": (c_int, [c_p, c_p, c_p, c_int, c_p]),
    "mdx_step_advance": (c_int, [c_p, c_int, c_p]),
    "mdx_gather_row": (c_int, [c_p, c_i64, c_p, c_int, c_p, c_int, c_int, c_p]),
    "mdx_sampler_mask_blend": (c_int, [c_p, c_p, c_p, c_int, c_f, c_f, c_p, c_int, c_int, c_int, c_int, c_p]),
    "mdx_sizeof_gemm_args": (c_int, []),
    "mdx_sizeof_conv_args": (c_int, []),
    "mdx_sizeof_step_args": (c_int, []),
    "mdx_offsetof_gemm_args_chan_stats": (c_int, []),
    "mdx_sampler_glide_step": (c_int, [c_p, c_i64, c_int, c_int, c_f, c_int, c_p, c_p, c_p, c_p, c_p, c_int, c_int,
                                       c_int, c_int, c_p]),
}

_lib = None


def lib_path():
    return _build.LIB


def load():
    """Build (if stale) and dlopen the kernel library. Raises if either step fails."""
    global _lib
    if _lib is not None:
        return _lib
    path = os.environ.get("MDX_LIB") or _build.build()   # MDX_LIB: a pre-built variant of the library (A/B runs)
    if not os.path.exists(path):
        raise MdxError("libmdx_b200.so is missing at %s; run `python -m minddiffusion_b200.build`" % path)
    lib = ctypes.CDLL(path)
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(lib, name)  # AttributeError if the .so does not export a declared symbol
        fn.restype = res
        fn.argtypes = args
    # the ctypes mirrors of the argument structs must match the compiled header byte for byte
    for name, cls in (("gemm", GemmArgs), ("conv", ConvArgs), ("step", StepArgs)):
        want = getattr(lib, "mdx_sizeof_%s_args" % name)()
        if ctypes.sizeof(cls) != want:
            raise MdxError("struct mdx_%s_args is %d bytes in the library, %d in the binding" %
                           (name, want, ctypes.sizeof(cls)))
    _lib = lib
    return lib


def check(rc, what):
    if rc != 0:
        msg = load().mdx_last_error()
        raise MdxError("%s failed (%d): %s" % (what, rc, msg.decode() if msg else "?"))


def launch_count():
    return int(load().mdx_launch_count())
