"""AutoencoderKL.decode on the sm_100a kernels (runs once per prompt batch, after the sampler loop).

Mirror of the decode path of the reference's first-stage model: `AutoencoderKL.decode`
(vision/stablediffusionv2/ldm/models/autoencoder.py:65-68) -> `post_quant_conv` -> `Decoder.construct`
(ldm/modules/diffusionmodules/model.py:408-440) with `ResnetBlock` (80-148), the single-head `AttnBlock` (151-206)
and `Upsample` (31-52).  Same constructor keywords as the YAML `ddconfig`, same parameter names
(`first_stage_model.decoder.up.3.block.0.norm1.gamma`, ...).  The encoder (training / inpainting only) is out of
scope.  Kernels: the UNet's GroupNorm+SiLU, tcgen05 implicit-GEMM conv (1x1 nin_shortcut K-concatenated behind the
second conv), nearest-2x upsample; the AttnBlock's 512-wide single head goes through GEMM -> row softmax -> GEMM.
"""
import numpy as np
import torch

from . import ops
from .unet import fold_upsample_weights, upsample_conv3x3

PREFIX = "first_stage_model."

SD_DDCONFIG = dict(double_z=True, z_channels=4, resolution=256, in_channels=3, out_ch=3, ch=128, ch_mult=[1, 2, 4, 4],
                   num_res_blocks=2, attn_resolutions=[], dropout=0.0)  # configs/v2-inference.yaml:44-60


def decoder_structure(dd):
    """Layer list of Decoder.__init__ (model.py:343-402), in execution order.
    ("conv_in", z, c) | ("res", name, cin, cout) | ("attn", name, c) | ("up", name, c) | ("out", c, out_ch)."""
    ch, mult, nrb = dd["ch"], list(dd["ch_mult"]), dd["num_res_blocks"]
    nres = len(mult)
    block_in = ch * mult[nres - 1]
    curr_res = dd["resolution"] // 2 ** (nres - 1)
    L = [("conv_in", dd["z_channels"], block_in), ("res", "mid.block_1", block_in, block_in),
         ("attn", "mid.attn_1", block_in), ("res", "mid.block_2", block_in, block_in)]
    for i_level in reversed(range(nres)):
        block_out = ch * mult[i_level]
        for i_block in range(nrb + 1):
            L.append(("res", "up.%d.block.%d" % (i_level, i_block), block_in, block_out))
            block_in = block_out
            if curr_res in dd["attn_resolutions"]:
                L.append(("attn", "up.%d.attn.%d" % (i_level, i_block), block_in))
        if i_level != 0:  # construct() skips the level-0 upsample (model.py:427-428) although the cell exists
            L.append(("up", "up.%d.upsample" % i_level, block_in))
            curr_res *= 2
    L.append(("out", block_in, dd["out_ch"]))
    return L


def decoder_param_shapes(dd, embed_dim=4, prefix=PREFIX):
    S = {prefix + "post_quant_conv.weight": (dd["z_channels"], embed_dim, 1, 1),
         prefix + "post_quant_conv.bias": (dd["z_channels"],)}
    d = prefix + "decoder."

    def conv(name, cout, cin, k):
        S[d + name + ".weight"] = (cout, cin, k, k)
        S[d + name + ".bias"] = (cout,)

    def norm(name, c):
        S[d + name + ".gamma"] = (c,)
        S[d + name + ".beta"] = (c,)

    for l in decoder_structure(dd):
        if l[0] == "conv_in":
            conv("conv_in", l[2], l[1], 3)
        elif l[0] == "res":
            _, nm, cin, cout = l
            norm(nm + ".norm1", cin)
            conv(nm + ".conv1", cout, cin, 3)
            norm(nm + ".norm2", cout)
            conv(nm + ".conv2", cout, cout, 3)
            if cin != cout:
                conv(nm + ".nin_shortcut", cout, cin, 1)
        elif l[0] == "attn":
            _, nm, c = l
            norm(nm + ".norm", c)
            for t in ("q", "k", "v", "proj_out"):
                conv(nm + "." + t, c, c, 1)
        elif l[0] == "up":
            conv(l[1] + ".conv", l[2], l[2], 3)
        elif l[0] == "out":
            norm("norm_out", l[1])
            conv("conv_out", l[2], l[1], 3)
    # the level-0 Upsample cell is constructed (and has parameters) but never called
    c0 = dd["ch"] * dd["ch_mult"][0]
    conv("up.0.upsample.conv", c0, c0, 3)
    return S


def random_decoder_params(dd, seed=0, embed_dim=4, prefix=PREFIX):
    rng = np.random.default_rng(seed)
    P = {}
    for name, shape in decoder_param_shapes(dd, embed_dim, prefix).items():
        if name.endswith(".gamma"):
            v = 1.0 + 0.1 * rng.standard_normal(shape)
        elif name.endswith(".beta") or name.endswith(".bias"):
            v = 0.05 * rng.standard_normal(shape)
        else:
            gain = 0.5 if (".conv2." in name or "proj_out" in name) else 1.0
            v = gain * rng.standard_normal(shape) / np.sqrt(int(np.prod(shape[1:])))
        P[name] = v.astype(np.float32)
    return P


class AutoencoderKL:
    """decode(z) only. z: fp32 NCHW latents [B, z_channels, h, w] -> fp32 NCHW image [B, out_ch, 8h, 8w]."""

    def __init__(self, params, ddconfig=None, embed_dim=4, dtype=torch.bfloat16, device="cuda", prefix=PREFIX):
        if not torch.cuda.is_available():
            raise RuntimeError("minddiffusion_b200 needs a CUDA device (sm_100a); there is no CPU path")
        self.dd = dict(ddconfig or SD_DDCONFIG)
        assert self.dd["double_z"]
        self.dtype, self.bf16 = dtype, dtype == torch.bfloat16
        self.device = torch.device(device)
        self.layers = decoder_structure(self.dd)
        self.zc, self.embed_dim = self.dd["z_channels"], embed_dim
        self.L = {}
        g = lambda n: self._get(params, prefix + n)
        w = lambda t: t.to(self.dtype).contiguous()
        ck = lambda t: t.permute(0, 2, 3, 1).reshape(t.shape[0], -1)
        self.L["pq"] = (g("post_quant_conv.weight").reshape(self.zc, embed_dim).contiguous(), g("post_quant_conv.bias"))
        d = "decoder."
        for l in self.layers:
            if l[0] == "conv_in":
                cw = g(d + "conv_in.weight")
                if self.bf16:  # pad z channels to 64 so conv_in runs on the tensor-core path
                    cp = torch.zeros((cw.shape[0], 64, 3, 3), device=self.device)
                    cp[:, :cw.shape[1]] = cw
                    cw = cp
                self.L["conv_in"] = (w(ck(cw)), g(d + "conv_in.bias"))
            elif l[0] == "res":
                _, nm, cin, cout = l
                self.L[nm + ".n1"] = (g(d + nm + ".norm1.gamma"), g(d + nm + ".norm1.beta"))
                self.L[nm + ".c1"] = (w(ck(g(d + nm + ".conv1.weight"))), g(d + nm + ".conv1.bias"))
                self.L[nm + ".n2"] = (g(d + nm + ".norm2.gamma"), g(d + nm + ".norm2.beta"))
                w2, b2 = ck(g(d + nm + ".conv2.weight")), g(d + nm + ".conv2.bias")
                if cin != cout:
                    ws, bs = g(d + nm + ".nin_shortcut.weight").reshape(cout, cin), g(d + nm + ".nin_shortcut.bias")
                    if self.bf16:
                        self.L[nm + ".c2"] = (w(torch.cat((w2, ws), 1)), (b2 + bs).contiguous())
                    else:
                        self.L[nm + ".c2"] = (w(w2), b2)
                        self.L[nm + ".nin"] = (w(ws), bs)
                else:
                    self.L[nm + ".c2"] = (w(w2), b2)
            elif l[0] == "attn":
                _, nm, c = l
                self.L[nm + ".n"] = (g(d + nm + ".norm.gamma"), g(d + nm + ".norm.beta"))
                for t in ("q", "k", "v", "proj_out"):
                    self.L[nm + "." + t] = (w(g(d + nm + "." + t + ".weight").reshape(c, c)), g(d + nm + "." + t + ".bias"))
            elif l[0] == "up":
                if self.bf16:  # nearest x2 + conv3x3 as four 2x2 phase GEMMs over the source image (unet.py)
                    self.L[l[1]] = ([w(wp) for wp in fold_upsample_weights(g(d + l[1] + ".conv.weight"))],
                                    g(d + l[1] + ".conv.bias"))
                else:
                    self.L[l[1]] = (w(ck(g(d + l[1] + ".conv.weight"))), g(d + l[1] + ".conv.bias"))
            elif l[0] == "out":
                self.L["norm_out"] = (g(d + "norm_out.gamma"), g(d + "norm_out.beta"))
                cw, cb = ck(g(d + "conv_out.weight")), g(d + "conv_out.bias")
                if self.bf16:  # pad the 3 output channels to 16 for the tensor-core path
                    wp = torch.zeros((16, cw.shape[1]), device=self.device)
                    wp[:cw.shape[0]] = cw
                    bp = torch.zeros(16, device=self.device)
                    bp[:cb.shape[0]] = cb
                    cw, cb = wp, bp
                self.L["conv_out"] = (w(cw), cb.contiguous())

    def _get(self, params, key):
        if key not in params:
            raise KeyError("missing parameter %s" % key)
        v = params[key]
        if isinstance(v, np.ndarray):
            v = torch.from_numpy(np.ascontiguousarray(v))
        return v.to(device=self.device, dtype=torch.float32).contiguous()

    # ---- op helpers ----
    def _chan(self, n, H, W, C, up=False):
        """Slot buffer for the GroupNorm statistics a GEMM epilogue collects for its consumer (see UNetModel._chan)."""
        slots = ops.stat_slots(n, H, W, up)
        return torch.empty((n, slots, 2, C), dtype=torch.float32, device=self.device) if slots else None

    def _conv3(self, x, w, b, residual=None, skips=None, trunk=False):
        n, H, W, _ = x.shape
        if self.bf16:
            segs = [(x, 9)] + [(s, 1) for s in (skips or [])]
            chan = self._chan(n, H, W, w.shape[0])
            out = ops.gemm_tc(segs, w, bias=b, residual=residual, chan_stats=chan,
                              out_dtype=torch.float32 if trunk else torch.bfloat16).view(n, H, W, -1)
            out._mdx_chan = chan
            return out
        return ops.conv_direct(x, w, 3, 3, 1, 1, bias=b, residual=residual)

    def _dense(self, x2, w, b=None, residual=None, trunk=False):
        if self.bf16:
            return ops.gemm_tc([(x2, 1)], w, bias=b, residual=residual,
                               out_dtype=torch.float32 if trunk else torch.bfloat16)
        return ops.conv_direct(x2, w, bias=b, residual=residual)

    def _res(self, nm, cin, cout, x):
        L, dt = self.L, self.dtype
        g1, b1 = L[nm + ".n1"]
        raw = None
        if self.bf16 and cin != cout:
            h, raw = ops.groupnorm(x, g1, b1, 32, 1e-6, silu=True, scratch=self._scr, out_dtype=dt, raw_copy=True,
                                   chan1=getattr(x, "_mdx_chan", None))
        else:
            h = ops.groupnorm(x, g1, b1, 32, 1e-6, silu=True, scratch=self._scr, out_dtype=dt,
                              chan1=getattr(x, "_mdx_chan", None))
        h = self._conv3(h, *L[nm + ".c1"])
        h = ops.groupnorm(h, *L[nm + ".n2"], 32, 1e-6, silu=True, scratch=self._scr, chan1=getattr(h, "_mdx_chan", None))
        w2, b2 = L[nm + ".c2"]
        if cin == cout:
            return self._conv3(h, w2, b2, residual=x, trunk=True)
        if self.bf16:
            return self._conv3(h, w2, b2, skips=[raw], trunk=True)
        sk = ops.conv_direct(x, *L[nm + ".nin"][:1], bias=L[nm + ".nin"][1])
        return self._conv3(h, w2, b2, residual=sk)

    def _attn(self, nm, c, x):
        """AttnBlock (model.py:186-206): softmax(q k^T c^-1/2) v with ONE head of width c."""
        L, dt = self.L, self.dtype
        n, H, W, _ = x.shape
        N = H * W
        hn = ops.groupnorm(x, *L[nm + ".n"], 32, 1e-6, silu=False, scratch=self._scr, out_dtype=dt).view(n * N, c)
        q = self._dense(hn, *L[nm + ".q"]).view(n, N, c)
        k = self._dense(hn, *L[nm + ".k"]).view(n, N, c)
        wv, bv = L[nm + ".v"]
        o = torch.empty((n, N, c), dtype=dt, device=x.device)
        for i in range(n):
            s = self._dense(q[i], k[i], trunk=True)                       # [N, N] scores, fp32
            p = ops.softmax_rows(s, float(c) ** -0.5, out_dtype=dt)
            vt = self._dense(wv, hn.view(n, N, c)[i])                     # V^T = Wv h^T : [c, N]
            # softmax rows sum to one, so the value bias passes through: P (V + 1 bv^T) = P V + bv^T
            if self.bf16:
                ops.gemm_tc([(p, 1)], vt, bias=bv, out=o[i])
            else:
                ops.conv_direct(p, vt, bias=bv, out=o[i])
        wp, bp = L[nm + ".proj_out"]
        return self._dense(o.view(n * N, c), wp, bp, residual=x.view(n * N, c), trunk=True).view(n, H, W, c)

    def decode(self, z):
        """autoencoder.py:65-68: post_quant_conv then the decoder."""
        L, dt = self.L, self.dtype
        z = z.to(device=self.device, dtype=torch.float32).contiguous()
        B, zc, h, w = z.shape
        self._scr = torch.empty(B * 64, dtype=torch.float64, device=self.device)
        z_nhwc = ops.nchw_to_nhwc(z, torch.float32)                       # [B, h, w, zc]
        cpad = 64 if self.bf16 else self.zc
        zq = torch.zeros((B, h, w, cpad), dtype=dt, device=self.device)
        ops.conv_direct(z_nhwc.view(B * h * w, zc), L["pq"][0], bias=L["pq"][1], out=zq.view(B * h * w, cpad)[:, :self.zc])
        x = None
        for l in self.layers:
            if l[0] == "conv_in":
                x = self._conv3(zq, *L["conv_in"], trunk=True)
            elif l[0] == "res":
                x = self._res(l[1], l[2], l[3], x)
            elif l[0] == "attn":
                x = self._attn(l[1], l[2], x)
            elif l[0] == "up":
                wu, bu = L[l[1]]
                if self.bf16:
                    chan = self._chan(x.shape[0], x.shape[1], x.shape[2], wu[0].shape[0], up=True)
                    x = upsample_conv3x3(ops.cast(x, dt), wu, bu, torch.float32, chan)
                    x._mdx_chan = chan
                else:
                    x = ops.conv_direct(x, wu, 3, 3, 1, 1, upsample=True, bias=bu)
            elif l[0] == "out":
                hh = ops.groupnorm(x, *L["norm_out"], 32, 1e-6, silu=True, scratch=self._scr, out_dtype=dt,
                                   chan1=getattr(x, "_mdx_chan", None))
                wo, bo = L["conv_out"]
                n, H, W, _ = hh.shape
                if self.bf16:
                    y = ops.gemm_tc([(hh, 9)], wo, bias=bo, out_dtype=torch.float32).view(n, H, W, -1)
                else:
                    y = ops.conv_direct(hh, wo, 3, 3, 1, 1, bias=bo)
                return ops.nhwc_to_nchw(y.contiguous(), self.dd["out_ch"])
        raise RuntimeError("decoder has no output layer")
