"""A torch-backed stand-in for the slice of the `mindspore` API that the reference's denoising path uses.

Purpose: MindSpore cannot be installed in this container, but the reference is pure Python on top of
it.  Installing this module as `mindspore` (see `install()`) lets `tests/golden/make_golden.py` import and EXECUTE
the UNMODIFIED reference sources under /root/reference (UNetModel, DiffusionWrapper, DDPM.register_schedule,
LatentDiffusion.apply_model, PLMSSampler, DPMSolverSampler, NoiseScheduleVP, ...) on CPU in fp32 and record golden
vectors for the oracle.  Every shim op maps one MindSpore primitive to the torch op with the same documented
semantics (MindSpore 1.9 docs): nn.Dense y = x W^T + b; nn.Conv2d pad_mode='pad'; nn.GroupNorm / nn.LayerNorm biased
variance, eps inside the sqrt; nn.GELU() / ops.GeLU tanh approximation; ResizeNearestNeighbor floor(dst * in/out);
type promotion of tensor-with-python-scalar keeps the tensor dtype (matters for the fp16 time grid of DPM-Solver).

This is test tooling only: nothing in the product imports it, and it runs only where /root/reference exists.
"""
import math
import sys
import types

import numpy as np
import torch
import torch.nn.functional as F

float16, float32, float64 = torch.float16, torch.float32, torch.float64
int32, int64, bool_ = torch.int32, torch.int64, torch.bool


class Tensor(torch.Tensor):
    """ms.Tensor(data, dtype=None) plus the MindSpore-only tensor methods the reference calls."""

    @staticmethod
    def __new__(cls, data=None, dtype=None, *args, **kwargs):
        if data is None and "input_data" in kwargs:
            data = kwargs.pop("input_data")
        if isinstance(data, torch.Tensor):
            t = data.detach().clone()
        elif isinstance(data, np.ndarray):
            t = torch.from_numpy(np.ascontiguousarray(data))
        elif isinstance(data, (list, tuple)) and len(data) and isinstance(data[0], torch.Tensor):
            t = torch.stack([torch.as_tensor(d) for d in data])
        else:
            t = torch.tensor(data)
            if t.dtype == torch.float64 and dtype is None:
                t = t.to(torch.float32)  # python floats become float32 tensors in MindSpore
        if dtype is not None:
            t = t.to(dtype)
        return torch.Tensor._make_subclass(cls, t, False)

    def asnumpy(self):
        return self.detach().cpu().numpy()

    def astype(self, dtype):
        return self.to(dtype)

    def expand_dims(self, axis):
        return self.unsqueeze(axis)

    def repeat(self, repeats, axis=None):  # numpy semantics (MindSpore), not torch.Tensor.repeat
        if axis is None:
            return torch.repeat_interleave(self.reshape(-1), repeats)
        return torch.repeat_interleave(self, repeats, dim=axis)

    def set_data(self, value):
        with torch.no_grad():
            self.copy_(value)
        return self

    def view(self, *shape):  # MindSpore's Tensor.view reshapes (copies when the layout requires it)
        if len(shape) == 1 and isinstance(shape[0], torch.dtype):
            return torch.Tensor.view(self, shape[0])
        return self.reshape(*shape)

    def transpose(self, *axes):  # MindSpore: a permutation (tuple or varargs), not torch's two-axis swap
        if len(axes) == 1 and isinstance(axes[0], (tuple, list)):
            axes = tuple(axes[0])
        return self.permute(*axes)


def tensor(data, dtype=None):
    return Tensor(data, dtype)


class Parameter(torch.nn.Parameter):
    def __new__(cls, data=None, requires_grad=True, name=None):
        return torch.nn.Parameter.__new__(cls, torch.as_tensor(data).detach().clone().float(), requires_grad)

    def set_data(self, value):
        with torch.no_grad():
            self.copy_(torch.as_tensor(value))
        return self


def _wrap(t):
    return t.as_subclass(Tensor) if isinstance(t, torch.Tensor) and not isinstance(t, Tensor) else t


# ---------------------------------------------------------------------------------------------------------
# mindspore.nn
# ---------------------------------------------------------------------------------------------------------
nn = types.ModuleType("mindspore.nn")


def _cast_floats(v, dtype):
    if isinstance(v, torch.Tensor) and v.is_floating_point() and v.dtype != dtype:
        return v.to(dtype)
    return v


def _set_to_float(module, dtype):
    """Cell.to_float: "add cast on all inputs of the cell and its child cells" (MindSpore docs)."""
    for m in module.modules():
        if isinstance(m, Cell):
            m._to_float = dtype
    return module


class Cell(torch.nn.Module):
    _to_float = None

    def forward(self, *args, **kwargs):
        if self._to_float is not None:
            args = tuple(_cast_floats(a, self._to_float) for a in args)
            kwargs = {k: _cast_floats(v, self._to_float) for k, v in kwargs.items()}
        return self.construct(*args, **kwargs)

    def to_float(self, dtype):
        return _set_to_float(self, dtype)

    param_prefix = ""

    def update_parameters_name(self, prefix="", recurse=True):  # torch names come from the attribute path already
        return None

    def set_train(self, mode=True):
        return self.train(mode)

    def parameters_and_names(self):
        return self.named_parameters()


class Dense(Cell):
    def __init__(self, in_channels, out_channels, weight_init="normal", bias_init="zeros", has_bias=True,
                 activation=None):
        super().__init__()
        self.weight = Parameter(torch.randn(out_channels, in_channels) * 0.01)
        self.bias = Parameter(torch.zeros(out_channels)) if has_bias else None

    def construct(self, x):
        return _wrap(F.linear(x, self.weight, self.bias))


class Conv2d(Cell):
    def __init__(self, in_channels, out_channels, kernel_size, stride=1, pad_mode="same", padding=0, dilation=1,
                 group=1, has_bias=False, weight_init="normal", bias_init="zeros", data_format="NCHW"):
        super().__init__()
        k = kernel_size if isinstance(kernel_size, (tuple, list)) else (kernel_size, kernel_size)
        self.stride, self.dilation, self.group = stride, dilation, group
        if pad_mode == "pad":
            self.padding = padding
        elif pad_mode == "valid":
            self.padding = 0
        elif pad_mode == "same":
            if stride != 1:
                raise NotImplementedError("shim: pad_mode='same' with stride != 1")
            self.padding = (k[0] // 2, k[1] // 2)
        else:
            raise ValueError(pad_mode)
        self.weight = Parameter(torch.randn(out_channels, in_channels // group, *k) * 0.01)
        self.bias = Parameter(torch.zeros(out_channels)) if has_bias else None

    def construct(self, x):
        return _wrap(F.conv2d(x, self.weight, self.bias, stride=self.stride, padding=self.padding,
                              dilation=self.dilation, groups=self.group))


class Conv1d(Cell):
    """nn.Conv1d keeps a 4-D weight [out, in, 1, k] (it runs on Conv2D internally)."""

    def __init__(self, in_channels, out_channels, kernel_size, stride=1, pad_mode="same", padding=0, dilation=1,
                 group=1, has_bias=False, weight_init="normal", bias_init="zeros"):
        super().__init__()
        if pad_mode not in ("pad", "valid") and kernel_size != 1:
            raise NotImplementedError("shim: Conv1d pad_mode=%r" % pad_mode)
        self.stride, self.padding = stride, (padding if pad_mode == "pad" else 0)
        self.weight = Parameter(torch.randn(out_channels, in_channels // group, 1, kernel_size) * 0.01)
        self.bias = Parameter(torch.zeros(out_channels)) if has_bias else None

    def construct(self, x):
        w = self.weight.reshape(self.weight.shape[0], self.weight.shape[1], -1)
        return _wrap(F.conv1d(x, w, self.bias, stride=self.stride, padding=self.padding))


class AvgPool2d(Cell):
    def __init__(self, kernel_size=1, stride=1, pad_mode="valid"):
        super().__init__()
        self.k, self.s = kernel_size, stride

    def construct(self, x):
        return _wrap(F.avg_pool2d(x, self.k, self.s))


class Embedding(Cell):
    def __init__(self, vocab_size, embedding_size, use_one_hot=False, embedding_table="normal", dtype=None,
                 padding_idx=None):
        super().__init__()
        self.embedding_table = Parameter(torch.randn(vocab_size, embedding_size) * 0.01)

    def construct(self, ids):
        return _wrap(self.embedding_table[ids.long()])


class _Linear(Cell):
    """mindspore.nn.transformer.layers._Linear: weight [out, in] (transpose_b=True), y = x W^T + b on the last axis."""

    def __init__(self, in_channels, out_channels, weight_init="normal", bias_init="zeros", has_bias=True,
                 activation=None, transpose_b=True, expert_num=1, outer_batch=1, expert_group_size=None,
                 param_init_type=None, compute_dtype=None):
        super().__init__()
        self.weight = Parameter(torch.randn(out_channels, in_channels) * 0.01)
        self.bias = Parameter(torch.zeros(out_channels))
        self.compute_dtype = compute_dtype

    def construct(self, x):
        dt = self.compute_dtype or x.dtype
        return _wrap(F.linear(x.to(dt), self.weight.to(dt), self.bias.to(dt)))


class BatchNorm2d(Cell):
    """Inference mode: moving statistics (the reference never puts the SRGAN generator in training mode)."""

    def __init__(self, num_features, eps=1e-5, momentum=0.9, affine=True, gamma_init="ones", beta_init="zeros",
                 moving_mean_init="zeros", moving_var_init="ones", use_batch_statistics=None, data_format="NCHW"):
        super().__init__()
        self.eps = eps
        self.gamma, self.beta = Parameter(torch.ones(num_features)), Parameter(torch.zeros(num_features))
        self.moving_mean = Parameter(torch.zeros(num_features), requires_grad=False)
        self.moving_variance = Parameter(torch.ones(num_features), requires_grad=False)

    def construct(self, x):
        return _wrap(F.batch_norm(x, self.moving_mean, self.moving_variance, self.gamma, self.beta, False, 0.0, self.eps))


class PReLU(Cell):
    def __init__(self, channel=1, w=0.25):
        super().__init__()
        self.w = Parameter(torch.full((channel,), float(w)))

    def construct(self, x):
        return _wrap(F.prelu(x, self.w))


class Tanh(Cell):
    def construct(self, x):
        return _wrap(torch.tanh(x))


class _Unused(Cell):
    def __init__(self, *a, **k):
        super().__init__()

    def construct(self, *a, **k):
        raise NotImplementedError("shim: layer not on the denoising path")


class GroupNorm(Cell):
    def __init__(self, num_groups, num_channels, eps=1e-5, affine=True, gamma_init="ones", beta_init="zeros"):
        super().__init__()
        self.num_groups, self.eps = num_groups, eps
        self.gamma = Parameter(torch.ones(num_channels))
        self.beta = Parameter(torch.zeros(num_channels))

    def construct(self, x):
        return _wrap(F.group_norm(x, self.num_groups, self.gamma, self.beta, self.eps))


class LayerNorm(Cell):
    def __init__(self, normalized_shape, begin_norm_axis=-1, begin_params_axis=-1, gamma_init="ones",
                 beta_init="zeros", epsilon=1e-7):
        super().__init__()
        self.shape = tuple(normalized_shape)
        self.eps = epsilon
        self.gamma = Parameter(torch.ones(self.shape))
        self.beta = Parameter(torch.zeros(self.shape))

    def construct(self, x):
        return _wrap(F.layer_norm(x, self.shape, self.gamma, self.beta, self.eps))


class SiLU(Cell):
    def construct(self, x):
        return _wrap(F.silu(x))


class GELU(Cell):
    def __init__(self, approximate=True):
        super().__init__()
        self.approximate = approximate

    def construct(self, x):
        return _wrap(F.gelu(x, approximate="tanh" if self.approximate else "none"))


class Dropout(Cell):
    def __init__(self, keep_prob=0.5, dtype=None):
        super().__init__()

    def construct(self, x):  # inference: identity
        return x


class SequentialCell(torch.nn.Sequential):
    def __init__(self, *args):
        if len(args) == 1 and isinstance(args[0], (list, tuple)):
            args = tuple(args[0])
        super().__init__(*args)

    def to_float(self, dtype):
        return _set_to_float(self, dtype)


class CellList(torch.nn.ModuleList):
    def __init__(self, modules=None, auto_prefix=True):
        super().__init__(modules)

    def to_float(self, dtype):
        return _set_to_float(self, dtype)


class MSELoss(Cell):
    def __init__(self, reduction="mean"):
        super().__init__()


class Sigmoid(Cell):
    def construct(self, x):
        return _wrap(torch.sigmoid(x))


class Softmax(Cell):
    def __init__(self, axis=-1):
        super().__init__()
        self.axis = axis

    def construct(self, x):
        return _wrap(torch.softmax(x, dim=self.axis))


for _n, _c in dict(Cell=Cell, Dense=Dense, Conv2d=Conv2d, Conv1d=Conv1d, Conv3d=_Unused, AvgPool1d=_Unused,
                   AvgPool2d=AvgPool2d, Embedding=Embedding, GroupNorm=GroupNorm, LayerNorm=LayerNorm, SiLU=SiLU, GELU=GELU,
                   Dropout=Dropout, BatchNorm2d=BatchNorm2d, PReLU=PReLU, Tanh=Tanh, Conv2dTranspose=_Unused,
                   SequentialCell=SequentialCell, Sequential=SequentialCell, CellList=CellList,
                   MSELoss=MSELoss, Softmax=Softmax, Sigmoid=Sigmoid, Pad=_Unused).items():
    setattr(nn, _n, _c)

# ---------------------------------------------------------------------------------------------------------
# mindspore.ops
# ---------------------------------------------------------------------------------------------------------
ops = types.ModuleType("mindspore.ops")


def _prim(fn):
    """A primitive class: instantiate (with attrs), then call."""

    class P:
        def __init__(self, *a, **k):
            self.a, self.k = a, k

        def __call__(self, *x, **kw):
            return _wrap(fn(self, *x, **kw))

        def add_prim_attr(self, *a, **k):
            return self

    return P


def _shape(s):
    return tuple(int(v) for v in s)


def _linspace(start, stop, num):
    start, stop = torch.as_tensor(start), torch.as_tensor(stop)
    return torch.linspace(float(start), float(stop), int(num), dtype=start.dtype if start.is_floating_point() else torch.float32)


def _split(x, axis=0, output_num=1):
    return tuple(_wrap(t) for t in torch.chunk(x, output_num, dim=axis))


def _broadcast_to(x, shape):
    return torch.as_tensor(x).expand(*_shape(shape)) if -1 not in shape else x.expand(*shape)


def _resize_bilinear_legacy(x, size):
    """ops.ResizeBilinear(size, align_corners=False) with its default half_pixel_centers=False: src = dst * in/out."""
    OH, OW = int(size[0]), int(size[1])
    H, W = x.shape[-2], x.shape[-1]
    fy = torch.arange(OH, dtype=torch.float32) * (np.float32(H) / np.float32(OH))
    fx = torch.arange(OW, dtype=torch.float32) * (np.float32(W) / np.float32(OW))
    y0, x0 = fy.floor().long(), fx.floor().long()
    y1, x1 = (y0 + 1).clamp(max=H - 1), (x0 + 1).clamp(max=W - 1)
    wy, wx = (fy - y0)[:, None], (fx - x0)[None, :]
    top = x[..., y0, :][..., x0] * (1 - wx) + x[..., y0, :][..., x1] * wx
    bot = x[..., y1, :][..., x0] * (1 - wx) + x[..., y1, :][..., x1] * wx
    return top * (1 - wy) + bot * wy


_fns = dict(
    bmm=torch.bmm, tanh=torch.tanh, transpose=lambda x, perm: x.permute(*perm),
    exp=torch.exp, expm1=torch.expm1, log=torch.log, sqrt=torch.sqrt, abs=torch.abs, sin=torch.sin, cos=torch.cos,
    pow=torch.pow, maximum=lambda a, b: torch.maximum(torch.as_tensor(a), torch.as_tensor(b)),
    minimum=lambda a, b: torch.minimum(torch.as_tensor(a), torch.as_tensor(b)),
    ones_like=torch.ones_like, ones=lambda shape, dtype=float32: torch.ones(_shape(shape), dtype=dtype),
    matmul=torch.matmul, reshape=lambda x, s: x.reshape(_shape(s)), expand_dims=lambda x, axis: x.unsqueeze(axis),
    concat=lambda xs, axis=0: torch.cat(list(xs), dim=axis), cast=lambda x, dt: x.to(dt),
    linspace=_linspace, broadcast_to=_broadcast_to, stop_gradient=lambda x: x.detach(),
    clip_by_value=lambda x, lo, hi: torch.max(torch.min(x, hi), lo),
    standard_normal=lambda shape, seed=0, seed2=0: torch.randn(_shape(shape)),
)
for _n, _f in _fns.items():
    setattr(ops, _n, (lambda f: (lambda *a, **k: _wrap(f(*a, **k))))(_f))
ops.split = _split


def _not_on_path(*a, **k):
    raise NotImplementedError("shim: op not on the denoising path")


ops.dropout = _not_on_path
ops.value_and_grad = _not_on_path

_prims = dict(
    Cast=lambda s, x, dt: torch.as_tensor(x).to(dt),
    Slice=lambda s, x, begin, size: x[tuple(slice(int(b), int(b) + int(n)) for b, n in zip(begin, size))],
    BroadcastTo=lambda s, x: torch.as_tensor(x).expand(*_shape(s.a[0])),
    Ones=lambda s, shape, dtype=float32: torch.ones(_shape(shape), dtype=dtype),
    Mul=lambda s, a, b: a * b, Add=lambda s, a, b: a + b, Div=lambda s, a, b: a / b, Neg=lambda s, x: -x,
    Round=lambda s, x: torch.round(x), Sin=lambda s, x: torch.sin(x), Print=lambda s, *a: None,
    UniformReal=lambda s, shape: torch.rand(_shape(shape)),
    ResizeBilinear=lambda s, x: _resize_bilinear_legacy(x, s.a[0]),
    DepthToSpace=lambda s, x: x.reshape(x.shape[0], s.a[0], s.a[0], x.shape[1] // (s.a[0] ** 2), x.shape[2],
                                        x.shape[3]).permute(0, 3, 4, 1, 5, 2).reshape(
        x.shape[0], x.shape[1] // (s.a[0] ** 2), x.shape[2] * s.a[0], x.shape[3] * s.a[0]),
    Zeros=lambda s, shape, dtype=float32: torch.zeros(_shape(shape), dtype=dtype),
    ZerosLike=lambda s, x: torch.zeros_like(x),
    Transpose=lambda s, x, perm: x.permute(*perm),
    Reshape=lambda s, x, shape: x.reshape(_shape(shape)),
    StandardNormal=lambda s, shape: torch.randn(_shape(shape)),
    Sqrt=lambda s, x: torch.sqrt(x), Square=lambda s, x: x * x, Log=lambda s, x: torch.log(x),
    Cos=lambda s, x: torch.cos(x), ACos=lambda s, x: torch.acos(x),
    Softmax=lambda s, x: torch.softmax(x, dim=s.a[0] if s.a else s.k.get("axis", -1)),
    Sigmoid=lambda s, x: torch.sigmoid(x),
    GeLU=lambda s, x: F.gelu(x, approximate="tanh"),
    Identity=lambda s, x: x,
    Exp=lambda s, x: torch.exp(x),
    Squeeze=lambda s, x: x.squeeze(s.a[0] if s.a else s.k.get("axis", ())),
    Gather=lambda s, params, idx, axis: torch.index_select(params, axis, idx.reshape(-1).long()).reshape(
        tuple(params.shape[:axis]) + tuple(idx.shape) + tuple(params.shape[axis + 1:])),
    AvgPool=lambda s, *a: (_ for _ in ()).throw(NotImplementedError("shim")),
    ExpandDims=lambda s, x, axis: x.unsqueeze(axis),
    Equal=lambda s, a, b: torch.eq(a, b),
    IsNan=lambda s, x: torch.isnan(x),
    GatherD=lambda s, x, dim, index: torch.gather(x, dim, index),
    ReverseV2=lambda s, x: torch.flip(x, dims=list(s.k.get("axis", s.a[0] if s.a else [0]))),
    Argmin=lambda s, x: torch.argmin(x.float(), dim=s.k.get("axis", s.a[0] if s.a else -1)),
    CumSum=lambda s, x, axis: torch.cumsum(x, dim=axis),
    LinSpace=lambda s, start, stop, num: _linspace(start, stop, num),
    Concat=lambda s, xs: torch.cat(list(xs), dim=s.a[0] if s.a else s.k.get("axis", 0)),
    ResizeNearestNeighbor=lambda s, x: F.interpolate(x, size=tuple(s.a[0]), mode="nearest"),
    UniformInt=lambda s, *a: (_ for _ in ()).throw(NotImplementedError("shim")),
    AvgPool3D=lambda s, *a: (_ for _ in ()).throw(NotImplementedError("shim")),
)
for _n, _f in _prims.items():
    setattr(ops, _n, _prim(_f))


class _SplitPrim:
    def __init__(self, axis=0, output_num=1):
        self.axis, self.n = axis, output_num

    def __call__(self, x):
        return _split(x, self.axis, self.n)


class _SortPrim:
    def __init__(self, axis=-1, descending=False):
        self.axis, self.desc = axis, descending

    def __call__(self, x):
        v, i = torch.sort(x, dim=self.axis, descending=self.desc, stable=True)
        return _wrap(v), _wrap(i)


ops.Split = _SplitPrim
ops.Sort = _SortPrim

# ---------------------------------------------------------------------------------------------------------
# mindspore.numpy, mindspore.common.initializer, top-level
# ---------------------------------------------------------------------------------------------------------
mnp = types.ModuleType("mindspore.numpy")


def _full(shape, fill_value, dtype=None):
    fv = torch.as_tensor(fill_value)
    if dtype is None:
        dtype = fv.dtype if fv.is_floating_point() or isinstance(fill_value, torch.Tensor) else (
            float32 if isinstance(fill_value, float) else int64)
        if dtype == torch.float64:
            dtype = float32
    return _wrap(torch.full(_shape(shape), fv.item(), dtype=dtype))


mnp.full = _full
mnp.where = lambda c, a, b: _wrap(torch.where(c.bool(), torch.as_tensor(a), torch.as_tensor(b)))
mnp.empty = lambda shape, dtype=float32: _wrap(torch.zeros(_shape(shape), dtype=dtype))
mnp.zeros = lambda shape, dtype=float32: _wrap(torch.zeros(_shape(shape), dtype=dtype))
mnp.ones = lambda shape, dtype=float32: _wrap(torch.ones(_shape(shape), dtype=dtype))
mnp.randint = lambda lo, hi=None, shape=(), dtype=int32: _wrap(torch.randint(lo, hi, _shape(shape)).to(dtype))
mnp.choose = lambda a, choices: _wrap(torch.as_tensor(choices)[a.long()])
mnp.tile = lambda x, reps: _wrap(torch.tile(x, tuple(reps) if isinstance(reps, (tuple, list)) else (reps,)))
mnp.flip = lambda x, axis=None: _wrap(torch.flip(x, dims=list(range(x.dim())) if axis is None else [axis]))
mnp.array = lambda data, dtype=None: Tensor(data, dtype)
mnp.arange = lambda start=0, stop=None, step=1, dtype=None: _wrap(
    torch.arange(start, stop, step, dtype=dtype) if stop is not None else torch.arange(start, dtype=dtype))
mnp.sqrt = lambda x: _wrap(torch.sqrt(x))
mnp.randn = lambda *shape: _wrap(torch.randn(*shape))

initializer_mod = types.ModuleType("mindspore.common.initializer")


class TruncatedNormal:
    def __init__(self, sigma=0.01):
        self.sigma = sigma


initializer_mod.TruncatedNormal = TruncatedNormal


class Normal:
    def __init__(self, sigma=0.01, mean=0.0):
        self.sigma, self.mean = sigma, mean


class XavierUniform:
    def __init__(self, gain=1):
        self.gain = gain


initializer_mod.Normal = Normal
initializer_mod.XavierUniform = XavierUniform


def initializer(init, shape=None, dtype=float32):
    if isinstance(init, TruncatedNormal):  # values are overwritten by the seeded weights anyway
        return Tensor(torch.randn(_shape(shape)).clamp_(-2, 2) * init.sigma, dtype)
    if isinstance(init, Normal):
        return Tensor(torch.randn(_shape(shape)) * init.sigma + init.mean, dtype)
    if init == "zeros":
        return Tensor(torch.zeros(_shape(shape), dtype=dtype))
    if init == "ones":
        return Tensor(torch.ones(_shape(shape), dtype=dtype))
    raise NotImplementedError("shim initializer %r" % (init,))


initializer_mod.initializer = initializer
common_mod = types.ModuleType("mindspore.common")
common_mod.initializer = initializer_mod


def install():
    """Register the shim as `mindspore` (and a stub `omegaconf`) in sys.modules. Returns the top-level module."""
    ms = types.ModuleType("mindspore")
    ms.__dict__.update(dict(Tensor=Tensor, tensor=tensor, Parameter=Parameter, float16=float16, float32=float32,
                            float64=float64, int32=int32, int64=int64, bool_=bool_, nn=nn, ops=ops, numpy=mnp,
                            common=common_mod, __version__="1.9-shim(torch)"))
    ms.dtype = types.SimpleNamespace(float16=float16, float32=float32, float64=float64, int32=int32, int64=int64,
                                     bool_=bool_)
    sys.modules["mindspore.dtype"] = ms.dtype
    ms.context = types.SimpleNamespace(set_context=lambda **k: None, GRAPH_MODE=0, PYNATIVE_MODE=1)
    ms.load_checkpoint = _not_on_path
    ms.load_param_into_net = _not_on_path
    sys.modules["mindspore"] = ms
    sys.modules["mindspore.nn"] = nn
    sys.modules["mindspore.ops"] = ops
    tr = types.ModuleType("mindspore.nn.transformer")
    trl = types.ModuleType("mindspore.nn.transformer.layers")
    trl._Linear = _Linear
    tr.layers = trl
    nn.transformer = tr
    sys.modules["mindspore.nn.transformer"] = tr
    sys.modules["mindspore.nn.transformer.layers"] = trl
    ops.operations = ops
    sys.modules["mindspore.ops.operations"] = ops
    sys.modules["mindspore.numpy"] = mnp
    sys.modules["mindspore.common"] = common_mod
    sys.modules["mindspore.common.initializer"] = initializer_mod
    oc = types.ModuleType("omegaconf")
    lc = types.ModuleType("omegaconf.listconfig")

    class ListConfig(list):
        pass

    lc.ListConfig = ListConfig
    oc.listconfig = lc
    oc.ListConfig = ListConfig
    sys.modules["omegaconf"] = oc
    sys.modules["omegaconf.listconfig"] = lc
    return ms
