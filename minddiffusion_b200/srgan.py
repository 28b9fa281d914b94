"""SRGAN x4 generator, the third stage of the Taichu-GLIDE cascade (256x256 -> 1024x1024), on the sm_100a kernels.

Host-side mirror of vision/Taichu-GLIDE/model/glide_text2im/model/srgan.py (`Generator`, `ResidualBlock`,
`SubpixelConvolutionLayer`) and srgan_util.py (`SRGAN(upscale_factor, ckpt).sr_handle(lr)`): same parameter names
(`conv1.0.weight`, `conv1.1.w`, `trunk.3.bn1.moving_mean`, `subpixel_conv.0.prelu.w`, ...), NCHW fp32 in
[-1, 1], NCHW fp32 out.  SURVEY.md §8f row 3.

  conv1   9x9 3->64 + PReLU                      CUDA-core implicit GEMM (Cin = 3 does not fill a tensor-core K block)
  trunk   16 x [conv3x3+BN -> PReLU -> conv3x3+BN -> + x]   BatchNorm (inference statistics) folded into the conv
                                                 weights at load; tcgen05 implicit GEMM, residual in the epilogue
  conv2   conv3x3 + PReLU, then conv1 + conv2    one PReLU kernel with the residual added after the activation
  2 x     conv3x3 64->256 -> DepthToSpace(2) -> PReLU       GEMM + one fused shuffle/PReLU mover
  conv3   9x9 64->3 + tanh                       CUDA-core kernel with the tanh epilogue

The reference runs this stage in fp32 (src/txt2img.py:129); dtype=torch.float32 does the same on the CUDA-core
kernels, dtype=torch.bfloat16 (default) keeps the residual trunk in fp32 and feeds the 3x3 convolutions bf16 operands.
"""
import numpy as np
import torch

from . import ops
from .unet import _t

BN_EPS = 1e-5  # nn.BatchNorm2d default


def srgan_param_shapes(upscale_factor=4):
    n_up = int(np.log2(upscale_factor))
    S = {"conv1.0.weight": (64, 3, 9, 9), "conv1.0.bias": (64,), "conv1.1.w": (64,),
         "conv2.0.weight": (64, 64, 3, 3), "conv2.0.bias": (64,), "conv2.1.w": (64,),
         "conv3.weight": (3, 64, 9, 9), "conv3.bias": (3,)}
    for i in range(16):
        p = "trunk.%d." % i
        for c in ("conv1", "conv2"):
            S[p + c + ".weight"] = (64, 64, 3, 3)
            S[p + c + ".bias"] = (64,)
        for b in ("bn1", "bn2"):
            for nm in ("gamma", "beta", "moving_mean", "moving_variance"):
                S[p + b + "." + nm] = (64,)
        S[p + "prelu.w"] = (64,)
    for i in range(n_up):
        p = "subpixel_conv.%d." % i
        S[p + "conv.weight"] = (256, 64, 3, 3)
        S[p + "conv.bias"] = (256,)
        S[p + "prelu.w"] = (64,)
    return S


def random_srgan_params(upscale_factor=4, seed=0):
    """Seeded stand-in weights (fan-in scaled; BatchNorm statistics and PReLU slopes away from their trivial values)."""
    rng = np.random.default_rng(seed)
    P = {}
    for name, shape in srgan_param_shapes(upscale_factor).items():
        if name.endswith("moving_variance"):
            v = 0.5 + rng.random(shape)
        elif name.endswith("gamma"):
            v = 1.0 + 0.1 * rng.standard_normal(shape)
        elif name.endswith((".beta", ".bias", "moving_mean")):
            v = 0.05 * rng.standard_normal(shape)
        elif name.endswith(".w"):
            v = 0.25 + 0.05 * rng.standard_normal(shape)
        else:
            gain = 0.5 if ".trunk." in "." + name and "conv2" in name else 1.0
            v = gain * rng.standard_normal(shape) / np.sqrt(np.prod(shape[1:]))
        P[name] = v.astype(np.float32)
    return P


class Generator:
    def __init__(self, params, upscale_factor=4, dtype=torch.bfloat16, device="cuda", prefix=""):
        if not torch.cuda.is_available():
            raise RuntimeError("minddiffusion_b200 needs a CUDA device (sm_100a); there is no CPU path")
        if dtype not in (torch.bfloat16, torch.float32):
            raise ValueError("dtype must be torch.bfloat16 or torch.float32")
        self.dtype, self.bf16 = dtype, dtype == torch.bfloat16
        self.device = torch.device(device)
        self.n_up = int(np.log2(upscale_factor))

        def g(name):
            key = prefix + name
            if key not in params:
                raise KeyError("missing parameter %s" % key)
            return _t(params[key], self.device)

        kmaj = lambda w: w.permute(0, 2, 3, 1).reshape(w.shape[0], -1).contiguous()
        wd = lambda w: kmaj(w).to(self.dtype).contiguous()

        def folded(conv, bn):  # conv followed by inference BatchNorm == conv with scaled weights and shifted bias
            s = g(bn + ".gamma") / torch.sqrt(g(bn + ".moving_variance") + BN_EPS)
            w = g(conv + ".weight") * s[:, None, None, None]
            b = (g(conv + ".bias") - g(bn + ".moving_mean")) * s + g(bn + ".beta")
            return wd(w), b.contiguous()

        self.c1 = (kmaj(g("conv1.0.weight")), g("conv1.0.bias"), g("conv1.1.w"))          # fp32: reads the fp32 image
        self.trunk = []
        for i in range(16):
            p = "trunk.%d." % i
            self.trunk.append((folded(p + "conv1", p + "bn1"), g(p + "prelu.w"), folded(p + "conv2", p + "bn2")))
        self.c2 = (wd(g("conv2.0.weight")), g("conv2.0.bias"), g("conv2.1.w"))
        self.up = [(wd(g("subpixel_conv.%d.conv.weight" % i)), g("subpixel_conv.%d.conv.bias" % i),
                    g("subpixel_conv.%d.prelu.w" % i)) for i in range(self.n_up)]
        self.c3 = (wd(g("conv3.weight")), g("conv3.bias"))

    def _conv3(self, x, w, b, residual=None, out_dtype=None):
        n, H, W, _ = x.shape
        if self.bf16:
            return ops.gemm_tc([(x, 9)], w, bias=b, residual=residual, out_dtype=out_dtype or torch.bfloat16).view(n, H, W, -1)
        return ops.conv_direct(x, w, 3, 3, 1, 1, bias=b, residual=residual)

    def __call__(self, x):
        """x: fp32 NCHW [n, 3, H, W] in [-1, 1] -> fp32 NCHW [n, 3, H*f, W*f] (Generator.construct, srgan.py:96-104)."""
        dt = self.dtype
        x = x.to(device=self.device, dtype=torch.float32).contiguous()
        n, _, H, W = x.shape
        h = ops.nchw_to_nhwc(x, torch.float32)
        w, b, a = self.c1
        c1 = ops.prelu(ops.conv_direct(h, w, 9, 9, 1, 4, bias=b), a)                       # fp32 [n, H, W, 64]
        t = c1                                                                             # fp32 residual trunk
        for (w1, b1), a1, (w2, b2) in self.trunk:
            u = self._conv3(ops.cast(t, dt) if self.bf16 else t, w1, b1)
            u = ops.prelu(u, a1)
            t = self._conv3(u, w2, b2, residual=t.view(n * H * W, 64), out_dtype=torch.float32)
        w, b, a = self.c2
        u = self._conv3(ops.cast(t, dt) if self.bf16 else t, w, b, out_dtype=torch.float32)
        t = ops.prelu(u, a, residual=c1, out_dtype=dt)                                     # conv1 + conv2
        for w, b, a in self.up:
            t = ops.depth_to_space2(self._conv3(t, w, b), a)
        w, b = self.c3
        y = ops.conv_direct(t, w, 9, 9, 1, 4, bias=b, act="tanh", out_dtype=torch.float32)
        return ops.nhwc_to_nchw(y.contiguous(), 3)


class SRGAN:
    """srgan_util.py:36-47: `SRGAN(upscale_factor, ckpt)`; weights are handed in as {name: array} instead of a
    checkpoint path (tools/ckpt_to_npz.py converts one)."""

    def __init__(self, upscale_factor=4, params=None, dtype=torch.bfloat16, device="cuda"):
        self.net = Generator(params if params is not None else random_srgan_params(upscale_factor), upscale_factor,
                             dtype=dtype, device=device)

    def sr_handle(self, lr):
        return self.net(lr)
