"""UNetModel of the latent-diffusion path, executed on the sm_100a kernels.

Host-side mirror of the reference's `UNetModel` (vision/stablediffusionv2/ldm/modules/diffusionmodules/
openaimodel.py:245-576; wukong-huahua's copy takes the SD-v1 constructor arguments): same constructor keywords
(the YAML `unet_config.params`), same parameter names (what `load_param_into_net` consumes), same call
`model(x, timesteps, context)`.  Underneath, activations are NHWC and every op is a kernel of libmdx_b200.so:

  ResBlock (openaimodel.py:176-205)   GroupNorm+SiLU pass -> conv3x3 (tcgen05 implicit GEMM, emb add in the
                                      epilogue) -> GroupNorm+SiLU -> conv3x3 with the 1x1 skip conv K-concatenated
                                      into the same accumulator (or the identity residual in the epilogue)
  SpatialTransformer (attention.py:187-256)  GroupNorm -> proj_in -> [LN -> fused QKV GEMM -> flash attention ->
                                      to_out(+residual)] -> [LN -> Q GEMM -> flash attention over the hoisted
                                      context K/V -> to_out(+residual)] -> [LN -> GEGLU GEMM -> Dense(+residual)]
                                      -> proj_out(+residual)
  skip concat (openaimodel.py:566-571) never copied: GroupNorm and the skip conv read both sources
  emb_layers of all ResBlocks         one small GEMM per forward (they depend only on t)

dtype=torch.bfloat16 is the product path; dtype=torch.float32 is the "check mode" of the north star: the same
graph on CUDA-core fp32 kernels, used to separate orchestration errors from bf16 rounding.
"""
import os

import numpy as np
import torch

from . import _lib, ops

PREFIX = "model.diffusion_model."


def unet_structure(cfg):
    """Block list implied by the constructor arguments (openaimodel.py:350-519).
    Layers: ("conv", cin, cout) | ("res", cin, cout) | ("st", ch, heads, dim_head) | ("down", ch) | ("up", ch)."""
    mc = cfg["model_channels"]
    state = {"heads": cfg.get("num_heads", -1)}
    nhc = cfg.get("num_head_channels", -1)
    legacy = cfg.get("legacy", True)
    if state["heads"] == -1 and nhc == -1:
        raise ValueError("Either num_heads or num_head_channels has to be set")  # openaimodel.py:317-321
    if not cfg.get("use_spatial_transformer", False):
        raise ValueError("only use_spatial_transformer=True is supported (AttentionBlock is an empty stub in the "
                         "reference, openaimodel.py:208-242)")
    if cfg.get("context_dim") is None:
        raise ValueError("context_dim must be set when use_spatial_transformer is True")  # openaimodel.py:305-306
    attn_res = [int(a) for a in cfg["attention_resolutions"]]
    mults = [int(m) for m in cfg["channel_mult"]]
    nrb = int(cfg["num_res_blocks"])

    def attn(ch):
        if nhc == -1:
            dh = ch // state["heads"]
        else:
            state["heads"] = ch // nhc
            dh = nhc
        if legacy:
            dh = ch // state["heads"]
        return ("st", ch, state["heads"], dh)

    inp = [[("conv", cfg["in_channels"], mc)]]
    chans = [mc]
    ch, ds = mc, 1
    for level, mult in enumerate(mults):
        for _ in range(nrb):
            block = [("res", ch, mult * mc)]
            ch = mult * mc
            if ds in attn_res:
                block.append(attn(ch))
            inp.append(block)
            chans.append(ch)
        if level != len(mults) - 1:
            inp.append([("down", ch)])
            chans.append(ch)
            ds *= 2
    mid = [("res", ch, ch), attn(ch), ("res", ch, ch)]
    out = []
    for level, mult in list(enumerate(mults))[::-1]:
        for i in range(nrb + 1):
            ich = chans.pop()
            block = [("res", ch + ich, mc * mult)]
            ch = mc * mult
            if ds in attn_res:
                block.append(attn(ch))
            if level and i == nrb:
                block.append(("up", ch))
                ds //= 2
            out.append(block)
    return inp, mid, out


def _t(v, device, dtype=torch.float32):
    if isinstance(v, np.ndarray):
        v = torch.from_numpy(np.ascontiguousarray(v))
    return v.to(device=device, dtype=dtype).contiguous()


class UNetModel:
    def __init__(self, params, dtype=torch.bfloat16, device="cuda", prefix=PREFIX, **cfg):
        if not torch.cuda.is_available():
            raise RuntimeError("minddiffusion_b200 needs a CUDA device (sm_100a); there is no CPU path")
        if dtype not in (torch.bfloat16, torch.float32):
            raise ValueError("dtype must be torch.bfloat16 (product) or torch.float32 (check mode)")
        self.cfg = dict(cfg)
        self.dtype = dtype
        self.bf16 = dtype == torch.bfloat16
        self.device = torch.device(device)
        self.prefix = prefix
        self.in_channels = cfg["in_channels"]
        self.out_channels = cfg["out_channels"]
        self.model_channels = cfg["model_channels"]
        self.context_dim = cfg["context_dim"]
        self.use_linear = cfg.get("use_linear_in_transformer", False)
        self.depth = cfg.get("transformer_depth", 1)
        self.inp, self.mid, self.out = unet_structure(cfg)
        self.cin_pad = 64 if self.bf16 else self.in_channels      # conv_in fed to the tensor-core path
        self.cout_pad = 16 if self.bf16 else self.out_channels     # conv_out likewise
        self.L = {}
        self._ctx = None
        self._concat = None          # 'concat' / 'hybrid' conditioning: extra input channels (set_concat)
        self._graphs = {}
        self.use_cuda_graph = False
        self.replayed_launches = 0  # kernels launched by graph replays (they bypass the C ABI's own counter)
        self._pack(params)

    # ------------------------------------------------------------------------------------------------
    # weight packing
    # ------------------------------------------------------------------------------------------------
    def _get(self, params, name):
        key = self.prefix + name
        if key not in params:
            raise KeyError("missing parameter %s" % key)
        return _t(params[key], self.device)

    def _w(self, w2d):
        return w2d.to(self.dtype).contiguous()

    @staticmethod
    def _conv_k(w):  # [Cout, Cin, kh, kw] -> [Cout, (r, s, c)]
        return w.permute(0, 2, 3, 1).reshape(w.shape[0], -1)

    def _pack(self, P):
        g = lambda n: self._get(P, n)
        L = self.L
        L["te0"] = (g("time_embed.0.weight"), g("time_embed.0.bias"))
        L["te2"] = (g("time_embed.2.weight"), g("time_embed.2.bias"))
        emb_w, emb_b, off = [], [], 0

        def pack_layer(pre, layer):
            nonlocal off
            kind = layer[0]
            if kind == "conv":
                w = g(pre + "conv.weight")
                if self.cin_pad != w.shape[1]:
                    wp = torch.zeros((w.shape[0], self.cin_pad, 3, 3), device=self.device)
                    wp[:, :w.shape[1]] = w
                    w = wp
                L[pre + "conv"] = (self._w(self._conv_k(w)), g(pre + "conv.bias"))
            elif kind == "res":
                cin, cout = layer[1], layer[2]
                L[pre + "n1"] = (g(pre + "in_layers_norm.gamma"), g(pre + "in_layers_norm.beta"))
                L[pre + "c1"] = (self._w(self._conv_k(g(pre + "in_layers_conv.conv.weight"))),
                                 g(pre + "in_layers_conv.conv.bias"))
                emb_w.append(g(pre + "emb_layers.1.weight"))
                emb_b.append(g(pre + "emb_layers.1.bias"))
                L[pre + "emb"] = (off, cout)
                off += cout
                L[pre + "n2"] = (g(pre + "out_layers_norm.gamma"), g(pre + "out_layers_norm.beta"))
                w2 = self._conv_k(g(pre + "out_layers_conv.conv.weight"))
                b2 = g(pre + "out_layers_conv.conv.bias")
                if cin != cout:
                    ws = g(pre + "skip_connection.conv.weight").reshape(cout, cin)
                    bs = g(pre + "skip_connection.conv.bias")
                    if self.bf16:  # K-concatenate the 1x1 skip conv behind the 3x3 taps
                        L[pre + "c2"] = (self._w(torch.cat((w2, ws), dim=1)), (b2 + bs).contiguous(), True)
                    else:
                        L[pre + "c2"] = (self._w(w2), b2, True)
                        L[pre + "skip"] = (self._w(ws), bs)
                else:
                    L[pre + "c2"] = (self._w(w2), b2, False)
            elif kind == "st":
                ch, heads, dh = layer[1], layer[2], layer[3]
                inner = heads * dh
                L[pre + "norm"] = (g(pre + "norm.gamma"), g(pre + "norm.beta"))
                L[pre + "proj_in"] = (self._w(g(pre + "proj_in.weight").reshape(inner, ch)), g(pre + "proj_in.bias"))
                L[pre + "proj_out"] = (self._w(g(pre + "proj_out.weight").reshape(ch, inner)), g(pre + "proj_out.bias"))
                for d in range(self.depth):
                    tb = pre + "transformer_blocks.%d." % d
                    for nm in ("norm1", "norm2", "norm3"):
                        L[tb + nm] = (g(tb + nm + ".gamma"), g(tb + nm + ".beta"))
                    L[tb + "qkv"] = self._w(torch.cat((g(tb + "attn1.to_q.weight"), g(tb + "attn1.to_k.weight"),
                                                       g(tb + "attn1.to_v.weight")), dim=0))
                    L[tb + "o1"] = (self._w(g(tb + "attn1.to_out.0.weight")), g(tb + "attn1.to_out.0.bias"))
                    L[tb + "q2"] = self._w(g(tb + "attn2.to_q.weight"))
                    L[tb + "kv2"] = self._w(torch.cat((g(tb + "attn2.to_k.weight"), g(tb + "attn2.to_v.weight")), dim=0))
                    L[tb + "o2"] = (self._w(g(tb + "attn2.to_out.0.weight")), g(tb + "attn2.to_out.0.bias"))
                    wg, bg = g(tb + "ff.net.0.proj.weight"), g(tb + "ff.net.0.proj.bias")
                    if self.bf16:
                        wg, bg, bn = pack_geglu(wg, bg)
                        L[tb + "ff1"] = (self._w(wg), bg.contiguous(), bn)
                    else:
                        L[tb + "ff1"] = (self._w(wg), bg, 0)
                    L[tb + "ff2"] = (self._w(g(tb + "ff.net.2.weight")), g(tb + "ff.net.2.bias"))
            elif kind == "down":
                L[pre + "op"] = (self._w(self._conv_k(g(pre + "op.conv.weight"))), g(pre + "op.conv.bias"))
            elif kind == "up":
                w = g(pre + "conv.conv.weight")
                if self.bf16:  # four 2x2 phase kernels instead of a 3x3 over the upsampled image
                    L[pre + "up"] = ([self._w(wp) for wp in fold_upsample_weights(w)], g(pre + "conv.conv.bias"))
                else:
                    L[pre + "up"] = (self._w(self._conv_k(w)), g(pre + "conv.conv.bias"))

        for bi, block in enumerate(self.inp):
            for li, layer in enumerate(block):
                pack_layer("input_blocks.%d.%d." % (bi, li), layer)
        for li, layer in enumerate(self.mid):
            pack_layer("middle_block.%d." % li, layer)
        for bi, block in enumerate(self.out):
            for li, layer in enumerate(block):
                pack_layer("output_blocks.%d.%d." % (bi, li), layer)
        L["emb_all"] = (torch.cat(emb_w, dim=0).contiguous(), torch.cat(emb_b, dim=0).contiguous())
        L["out_norm"] = (g("out.0.gamma"), g("out.0.beta"))
        w = self._conv_k(g("out.2.conv.weight"))
        b = g("out.2.conv.bias")
        if self.cout_pad != w.shape[0]:
            wp = torch.zeros((self.cout_pad, w.shape[1]), device=self.device)
            wp[:w.shape[0]] = w
            bp = torch.zeros(self.cout_pad, device=self.device)
            bp[:b.shape[0]] = b
            w, b = wp, bp
        L["out_conv"] = (self._w(w), b.contiguous())

    # ------------------------------------------------------------------------------------------------
    # op helpers (bf16 tensor-core path / fp32 check path)
    # ------------------------------------------------------------------------------------------------
    def _chan(self, n, H, W, C, up=False):
        """[n, slots, 2, C] fp32 buffer a GEMM epilogue fills with partial per-channel sums / sums of squares of its
        output (one slot per 32 rows): the GroupNorm that reads the tensor next needs no statistics pass
        (mdx_groupnorm_pre).  None for images below 128 pixels (two-pass GroupNorm there)."""
        if os.environ.get("MDX_NO_EPI_STATS"):   # A/B switch: two-pass GroupNorm everywhere
            return None
        slots = ops.stat_slots(n, H, W, up)
        return torch.empty((n, slots, 2, C), dtype=torch.float32, device=self.device) if slots else None

    def _conv3(self, x, w, b, rowbias=None, residual=None, skips=None, trunk=False, stats=False):
        """3x3 stride-1 conv on NHWC x; `skips`: raw bf16 tensors whose fused 1x1 conv is K-concatenated (bf16
        mode); trunk=True writes the fp32 residual stream; stats=True tags the result with its channel statistics."""
        n, H, W, _ = x.shape
        if self.bf16:
            segs = [(x, 9)] + [(s, 1) for s in (skips or [])]
            chan = self._chan(n, H, W, w.shape[0]) if stats else None
            out = ops.gemm_tc(segs, w, bias=b, rowbias=rowbias, residual=residual, chan_stats=chan,
                              out_dtype=torch.float32 if trunk else torch.bfloat16).view(n, H, W, -1)
            out._mdx_chan = chan
            return out
        return ops.conv_direct(x, w, 3, 3, 1, 1, bias=b, rowbias=rowbias, residual=residual)

    def _linear(self, x, w, b=None, residual=None, geglu_bn=None, out=None, trunk=False):
        """Dense layer on rows; x: [..., K] view. trunk=True writes the fp32 residual stream (bf16 mode)."""
        if self.bf16:
            x2 = x.reshape(-1, x.shape[-1]) if x.is_contiguous() else x
            if x2.dim() != 2:
                raise ValueError("non-contiguous dense input must be 2-D")
            r2 = residual.reshape(-1, residual.shape[-1]) if residual is not None else None
            if geglu_bn:
                return ops.gemm_tc([(x2, 1)], w, bias=b, geglu=True, block_n=geglu_bn, out=out)
            return ops.gemm_tc([(x2, 1)], w, bias=b, residual=r2, out=out,
                               out_dtype=torch.float32 if trunk else torch.bfloat16)
        x2 = x.reshape(-1, x.shape[-1])
        r2 = residual.reshape(-1, residual.shape[-1]) if residual is not None else None
        return ops.conv_direct(x2, w, bias=b, residual=r2, out=out)

    # ------------------------------------------------------------------------------------------------
    # blocks.  bf16 mode keeps the residual stream ("trunk": block inputs/outputs, the skip list and the
    # transformer's token stream) in fp32 and every GEMM operand in bf16; see DESIGN.md "precision".
    # ------------------------------------------------------------------------------------------------
    def _res(self, pre, layer, x1, x2, emb_all):
        L = self.L
        cin, cout = layer[1], layer[2]
        g1, b1 = L[pre + "n1"]
        w2, bb2, has_skip = L[pre + "c2"]
        raw = None
        ch1 = getattr(x1, "_mdx_chan", None)
        ch2 = getattr(x2, "_mdx_chan", None) if x2 is not None else None
        if self.bf16 and has_skip:  # the same pass also emits the bf16 operand of the 1x1 skip conv
            h, raw = ops.groupnorm(x1, g1, b1, 32, 1e-5, silu=True, x2=x2, scratch=self._gn_scratch,
                                   out_dtype=self.dtype, raw_copy=True, chan1=ch1, chan2=ch2)
        else:
            h = ops.groupnorm(x1, g1, b1, 32, 1e-5, silu=True, x2=x2, scratch=self._gn_scratch, out_dtype=self.dtype,
                              chan1=ch1, chan2=ch2)
        w1, bb1 = L[pre + "c1"]
        off, _ = L[pre + "emb"]
        h = self._conv3(h, w1, bb1, rowbias=emb_all[:, off:off + cout], stats=True)
        g2, b2 = L[pre + "n2"]
        h = ops.groupnorm(h, g2, b2, 32, 1e-5, silu=True, scratch=self._gn_scratch, chan1=getattr(h, "_mdx_chan", None))
        if not has_skip:
            return self._conv3(h, w2, bb2, residual=x1, trunk=True, stats=True)
        if self.bf16:
            return self._conv3(h, w2, bb2, skips=[raw], trunk=True, stats=True)
        xin = x1 if x2 is None else ops.concat_channels(x1, x2)
        ws, bs = L[pre + "skip"]
        sk = ops.conv_direct(xin, ws, bias=bs)
        return self._conv3(h, w2, bb2, residual=sk)

    def _st(self, pre, layer, x):
        L = self.L
        ch, heads, dh = layer[1], layer[2], layer[3]
        n, H, W, _ = x.shape
        N = H * W
        scale = dh ** -0.5
        dt = self.dtype
        gn, bn_ = L[pre + "norm"]
        hn = ops.groupnorm(x, gn, bn_, 32, 1e-6, silu=False, scratch=self._gn_scratch, out_dtype=dt,
                           chan1=getattr(x, "_mdx_chan", None))
        w, b = L[pre + "proj_in"]
        t = self._linear(hn, w, b, trunk=True)           # [n*N, C] token stream
        C = t.shape[-1]
        for d in range(self.depth):
            tb = pre + "transformer_blocks.%d." % d
            a = ops.layernorm(t, *L[tb + "norm1"], eps=1e-5, out_dtype=dt)
            qkv = self._linear(a, L[tb + "qkv"]).view(n, N, 3 * C)
            o = ops.attention(qkv[:, :, :C], qkv[:, :, C:2 * C], qkv[:, :, 2 * C:], heads, scale)
            t = self._linear(o, *L[tb + "o1"], residual=t, trunk=True)
            a = ops.layernorm(t, *L[tb + "norm2"], eps=1e-5, out_dtype=dt)
            q = self._linear(a, L[tb + "q2"]).view(n, N, C)
            kv = self._ctx[tb]
            o = ops.attention(q, kv[:, :, :C], kv[:, :, C:], heads, scale)
            t = self._linear(o, *L[tb + "o2"], residual=t, trunk=True)
            a = ops.layernorm(t, *L[tb + "norm3"], eps=1e-5, out_dtype=dt)
            wg, bg, gbn = L[tb + "ff1"]
            if self.bf16:
                gg = self._linear(a, wg, bg, geglu_bn=gbn)
            else:
                gg = ops.geglu(self._linear(a, wg, bg))
            # the last block's output only feeds proj_out as a GEMM operand: store it in the operand dtype
            t = self._linear(gg, *L[tb + "ff2"], residual=t, trunk=(d != self.depth - 1))
        w, b = L[pre + "proj_out"]
        if self.bf16:  # as a 1x1 conv over [n, H, W, C] so that the epilogue knows each row's image
            chan = self._chan(n, H, W, ch)
            out = ops.gemm_tc([(t.view(n, H, W, C), 1)], w, bias=b, residual=x.view(n * N, ch), chan_stats=chan,
                              out_dtype=torch.float32).view(n, H, W, ch)
            out._mdx_chan = chan
            return out
        return self._linear(t, w, b, residual=x, trunk=True).view(n, H, W, ch)

    def _down(self, pre, x):
        w, b = self.L[pre + "op"]
        n, H, W, C = x.shape
        if self.bf16:   # the nine taps are strided TMA boxes over the bf16 copy of the trunk: no patch matrix
            OH, OW = (H - 1) // 2 + 1, (W - 1) // 2 + 1
            chan = self._chan(n, OH, OW, w.shape[0])
            out = ops.gemm_tc([(ops.cast(x, self.dtype), 9)], w, bias=b, out_dtype=torch.float32, chan_stats=chan,
                              stride=2).view(n, OH, OW, -1)
            out._mdx_chan = chan
            return out
        return ops.conv_direct(x, w, 3, 3, 2, 1, bias=b)

    def _up(self, pre, x):
        w, b = self.L[pre + "up"]
        if self.bf16:
            chan = self._chan(x.shape[0], x.shape[1], x.shape[2], w[0].shape[0], up=True)
            out = upsample_conv3x3(ops.cast(x, self.dtype), w, b, torch.float32, chan)
            out._mdx_chan = chan
            return out
        return ops.conv_direct(x, w, 3, 3, 1, 1, upsample=True, bias=b)

    # ------------------------------------------------------------------------------------------------
    # public API
    # ------------------------------------------------------------------------------------------------
    def graph_pool(self):
        """Memory pool shared by the captured step graphs of samplers.py: they replay one after the other and keep
        nothing inside the pool between replays, so their temporaries may overlap."""
        if getattr(self, "_pool", None) is None:
            self._pool = torch.cuda.graph_pool_handle()
        return self._pool

    def set_concat(self, c_concat):
        """Channel-concatenated conditioning of the 'concat' / 'hybrid' DiffusionWrapper (wukong-huahua ddpm.py:363-372;
        inpainting: mask + masked-image latents, inpaint.py:78-90).  c_concat: [n, Cc, h, w] for the n images of a
        forward (n = B*repeat, e.g. cat((uc_concat, c_concat))) or [B, Cc, h, w] to share it between the repeats; None
        clears it.  The UNet's in_channels must equal the latent channels + Cc."""
        if c_concat is None:
            if self._concat is not None:
                self._concat = None
                self._graphs = {}  # captured graphs read the previous buffer
            return
        c = c_concat.to(device=self.device, dtype=torch.float32)
        if self._concat is not None and self._concat.shape == c.shape:
            self._concat.copy_(c)  # same shape: refill the buffer the captured graphs read
        else:
            self._concat = c.contiguous().clone()
            self._graphs = {}

    def set_context(self, context):
        """Hoist the cross-attention K/V projections (attention.py:118-121) out of the step loop: the context is
        constant across sampler steps.  context: [n, L, context_dim] (any float dtype, on the device)."""
        ctx = context.to(device=self.device, dtype=self.dtype).contiguous()
        n, Lc, D = ctx.shape
        if D != self.context_dim:
            raise ValueError("context dim %d != context_dim %d" % (D, self.context_dim))
        old = self._ctx if (self._ctx is not None and self._ctx["n"] == n and self._ctx["L"] == Lc) else None
        if old is None:
            self._ctx = {"n": n, "L": Lc}
            self._graphs = {}  # captured graphs point at the previous K/V buffers

        def visit(pre, layer):
            if layer[0] != "st":
                return
            for d in range(self.depth):
                tb = pre + "transformer_blocks.%d." % d
                if old is not None:  # same shapes: refill the buffers the captured graphs read
                    self._linear(ctx.view(n * Lc, D), self.L[tb + "kv2"], out=old[tb].view(n * Lc, -1))
                else:
                    kv = self._linear(ctx.view(n * Lc, D), self.L[tb + "kv2"])
                    self._ctx[tb] = kv.view(n, Lc, -1)

        for bi, block in enumerate(self.inp):
            for li, layer in enumerate(block):
                visit("input_blocks.%d.%d." % (bi, li), layer)
        for li, layer in enumerate(self.mid):
            visit("middle_block.%d." % li, layer)
        for bi, block in enumerate(self.out):
            for li, layer in enumerate(block):
                visit("output_blocks.%d.%d." % (bi, li), layer)

    def emb_rows(self, t):
        """timestep embedding -> time_embed MLP -> every ResBlock's emb_layers (openaimodel.py:548-549, 199-200), all
        fp32: t [m] -> [m, sum of the ResBlocks' Cout].  The result depends on nothing but t, so a sampler loop
        computes the rows of all its steps in one call before the loop (samplers._StepPlan)."""
        L = self.L
        t_emb = ops.timestep_embedding(t, self.model_channels)
        emb = ops.conv_direct(t_emb, L["te0"][0], bias=L["te0"][1])
        emb = ops.conv_direct(emb, L["te2"][0], bias=L["te2"][1], pre_silu=True)
        return ops.conv_direct(emb, L["emb_all"][0], bias=L["emb_all"][1], pre_silu=True)

    def _forward_impl(self, x_nchw, t, repeat, taps=None, uniform_t=False, emb_row=None):
        L = self.L
        n = x_nchw.shape[0] * repeat
        self._gn_scratch = torch.empty(n * 64, dtype=torch.float64, device=self.device)
        # The sampler loops evaluate the whole batch at ONE timestep (plms.py:118, ddim ts = full((b,), step)): one
        # embedding row is computed (or handed in: emb_row [1, sum Cout]) and every image reads it (row stride 0 of the
        # per-image bias).
        if emb_row is not None:
            emb_all = emb_row.expand(n, -1)
        else:
            emb_all = self.emb_rows(t[:1] if uniform_t else t)  # [n, sum Cout]
            if uniform_t:
                emb_all = emb_all.expand(n, -1)
        if self._concat is not None:   # DiffusionWrapper 'concat' / 'hybrid' (wukong ddpm.py:363-372): cat((x, c_concat), 1)
            h = ops.nchw2_to_nhwc(x_nchw, self._concat, self.dtype, repeat=repeat, cpad=self.cin_pad)
        else:
            h = ops.nchw_to_nhwc(x_nchw, self.dtype, repeat=repeat, cpad=self.cin_pad)
        hs = []

        def run(pre, layer, h, skip):
            kind = layer[0]
            if kind == "conv":
                return self._conv3(h, *L[pre + "conv"], trunk=True, stats=True)
            if kind == "res":
                return self._res(pre, layer, h, skip, emb_all)
            if kind == "st":
                return self._st(pre, layer, h)
            if kind == "down":
                return self._down(pre, h)
            return self._up(pre, h)

        for bi, block in enumerate(self.inp):
            for li, layer in enumerate(block):
                h = run("input_blocks.%d.%d." % (bi, li), layer, h, None)
            hs.append(h)
            if taps is not None:
                taps["input_blocks.%d" % bi] = h
        for li, layer in enumerate(self.mid):
            h = run("middle_block.%d." % li, layer, h, None)
        if taps is not None:
            taps["middle_block"] = h
        for bi, block in enumerate(self.out):
            skip = hs.pop()
            for li, layer in enumerate(block):
                h = run("output_blocks.%d.%d." % (bi, li), layer, h, skip if li == 0 else None)
            if taps is not None:
                taps["output_blocks.%d" % bi] = h
        g, b = L["out_norm"]
        h = ops.groupnorm(h, g, b, 32, 1e-5, silu=True, scratch=self._gn_scratch, out_dtype=self.dtype,
                          chan1=getattr(h, "_mdx_chan", None))
        nn_, H, W, _ = h.shape
        w, b = L["out_conv"]
        if self.bf16:
            return ops.gemm_tc([(h, 9)], w, bias=b, out_dtype=torch.float32).view(nn_, H, W, -1)
        return ops.conv_direct(h, w, 3, 3, 1, 1, bias=b)

    def forward_nhwc(self, x_nchw, timesteps, repeat=1, taps=None, uniform_t=False):
        """x_nchw: fp32 [B, C, H, W]; timesteps: [B*repeat] (any numeric dtype); the batch is read `repeat` times
        (repeat=2 is the CFG concat of plms.py:192-194 without the copy).  Returns fp32 NHWC [B*repeat, H, W, ld]
        with ld >= out_channels (bf16 mode pads conv_out to 16 columns)."""
        if self._ctx is None:
            raise RuntimeError("call set_context(context) before forward")
        n = x_nchw.shape[0] * repeat
        if self._ctx["n"] != n:
            raise ValueError("context batch %d != input batch %d" % (self._ctx["n"], n))
        x_nchw = x_nchw.to(device=self.device, dtype=torch.float32).contiguous()
        t = timesteps.to(device=self.device, dtype=torch.float32).contiguous()
        if t.numel() != n:
            raise ValueError("timesteps must have %d entries" % n)
        cc = 0 if self._concat is None else self._concat.shape[1]
        if x_nchw.shape[1] + cc != self.in_channels:
            raise ValueError("input has %d (+%d concat) channels, the UNet takes %d" % (x_nchw.shape[1], cc, self.in_channels))
        if self._concat is not None and (self._concat.shape[0] not in (x_nchw.shape[0], n)
                                         or self._concat.shape[2:] != x_nchw.shape[2:]):
            raise ValueError("c_concat must be [B or B*repeat, Cc, h, w] matching the input")
        if not self.use_cuda_graph or taps is not None:
            return self._forward_impl(x_nchw, t, repeat, taps, uniform_t)
        key = (tuple(x_nchw.shape), repeat, uniform_t)
        ent = self._graphs.get(key)
        if ent is None:
            sx, st_ = x_nchw.clone(), t.clone()
            self._forward_impl(sx, st_, repeat, None, uniform_t)  # warm-up: sets kernel attributes, fills the allocator
            torch.cuda.synchronize()
            graph = torch.cuda.CUDAGraph()
            n0 = _lib.launch_count()
            with torch.cuda.graph(graph):
                sout = self._forward_impl(sx, st_, repeat, None, uniform_t)
            ent = (graph, sx, st_, sout, _lib.launch_count() - n0)
            self._graphs[key] = ent
        graph, sx, st_, sout, nk = ent
        sx.copy_(x_nchw)
        st_.copy_(t)
        graph.replay()
        self.replayed_launches += nk
        return sout

    def __call__(self, x, timesteps=None, context=None, y=None):
        """Reference call signature (openaimodel.py:536): NCHW in, NCHW eps out."""
        if y is not None:
            raise ValueError("class-conditional UNet (num_classes) is not used by any shipped config")
        if context is not None:
            self.set_context(context)
        out = self.forward_nhwc(x, timesteps)
        return ops.nhwc_to_nchw(out.contiguous(), self.out_channels)


def fold_upsample_weights(w):
    """Upsample = nearest x2 then conv3x3 (openaimodel.py:53-63).  Output pixel (2y+py, 2x+px) reads upsampled rows
    2y+py+r, r in {-1, 0, 1}, i.e. source rows floor((2y+py+r)/2): for py = 0 {y-1 | y, y}, for py = 1 {y, y | y+1}.
    So each output phase is a 2x2 convolution of the SOURCE image whose taps are sums of the 3x3 taps that land on
    the same source pixel: 4/9 of the multiply-adds and no upsampled operand.  Returns the four phase matrices
    [Cout, (a, b, c)] in phase order (py, px) = (0,0), (0,1), (1,0), (1,1); window offset is (py-1, px-1)."""
    groups = {0: ((0,), (1, 2)), 1: ((0, 1), (2,))}   # phase -> kernel rows (or columns) folded into window tap 0 / 1
    out = []
    for py in (0, 1):
        for px in (0, 1):
            taps = []
            for a in (0, 1):
                for b_ in (0, 1):
                    acc = 0
                    for r in groups[py][a]:
                        for c in groups[px][b_]:
                            acc = acc + w[:, :, r, c]
                    taps.append(acc)                       # [Cout, Cin]
            out.append(torch.stack(taps, dim=1).reshape(w.shape[0], -1).contiguous())
    return out


def upsample_conv3x3(x, phase_weights, bias, out_dtype, chan_stats=None):
    """x: bf16 NHWC [n, H, W, C] -> conv3x3(nearest_x2(x)) as [n, 2H, 2W, Cout], four mdx_gemm_tc launches."""
    n, H, W, _ = x.shape
    out = torch.empty((n, 2 * H, 2 * W, phase_weights[0].shape[0]), dtype=out_dtype, device=x.device)
    for ph, wp in enumerate(phase_weights):
        ops.gemm_tc([(x, 4)], wp, out=out.view(n * 4 * H * W, -1), bias=bias, up_phase=ph + 1, chan_stats=chan_stats)
    return out


def pack_geglu(w, b):
    """Re-order GEGLU.proj rows (attention.py:44: [value rows | gate rows]) so that every BLOCK_N-wide output tile
    of the GEMM holds [BLOCK_N/2 value rows | the matching BLOCK_N/2 gate rows]; the epilogue then multiplies
    value * gelu(gate) without another pass (include/mdx.h, mdx_gemm_tc geglu)."""
    two_i = w.shape[0]
    inner = two_i // 2
    bn = ops.pick_block_n(two_i, geglu=True)
    if bn == 0:
        raise ValueError("GEGLU width %d has no valid tile" % two_i)
    half = bn // 2
    idx = []
    for t in range(inner // half):
        idx.extend(range(t * half, (t + 1) * half))
        idx.extend(range(inner + t * half, inner + (t + 1) * half))
    idx = torch.tensor(idx, device=w.device)
    return w[idx].contiguous(), b[idx].contiguous(), bn
