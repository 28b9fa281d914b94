"""Taichu-GLIDE on the sm_100a kernels: text-conditioned pixel-space UNet (64x64 base, 256x256 upsampler), its
per-step text transformer, and the learned-sigma samplers.

Host-side mirror of vision/Taichu-GLIDE/model/glide_text2im:
  default_options.py     model_and_diffusion_defaults / model_and_diffusion_upsample
  gaussian_computation.py, diffusion_creator.py:86-95   beta schedules, respacing ("60", "fast27", "ddimN")
  model/unet.py          ResBlock (scale-shift norm, resblock up/down), AttentionBlock with encoder_kv, UNetModel
  model/text2im_model.py Text2ImUNet / SuperResText2ImUNet (token + positional embedding, xf_padding, transformer,
                         final_ln, transformer_proj added to the time embedding, xf_out as encoder K/V source)
  model/xf.py            the text transformer
  model/guider.py, model/gaussian_diffusion.py   CombinePrompt + Guider + PMeanVariance + PSample / DDimSample
  main_funcs.py          gaussian_p_sample_loop, ddim_sample_loop
with the same option keys, parameter names and call signatures.  Activations are NHWC; every op is a kernel of
libmdx_b200.so.  Work the reference repeats every step but that does not change between steps is hoisted:
  * the prompt half's text transformer, its transformer_proj row and its encoder K/V run once per prompt;
  * the random-prompt half (main_funcs.py:33, one random row broadcast to all images, guider.py:36-41) runs the text
    transformer on ONE row per step instead of `pics_generated` identical rows;
  * the upsampler's low-res conditioning (quantise + bilinear resize, gaussian_diffusion.py:224-230,
    text2im_model.py:191) is computed once per sample.
dtype=torch.bfloat16 is the product path, torch.float32 the check mode (same graph on CUDA-core fp32 kernels).
"""
import math

import numpy as np
import torch

from . import _lib, ops
from .unet import _t, fold_upsample_weights, upsample_conv3x3


# -------------------------------------------------------------------------------------------------------------
# options (default_options.py:18-148)
# -------------------------------------------------------------------------------------------------------------
def model_and_diffusion_defaults(**over):
    o = dict(image_size=64, num_channels=192, num_res_blocks=3, channel_mult=(1, 2, 3, 4), num_heads=1,
             num_head_channels=64, num_heads_upsample=-1, attention_resolutions=(2, 4, 8), dropout=0.9, text_ctx=128,
             xf_width=512, xf_layers=16, xf_heads=8, xf_final_ln=True, n_vocab=50001, xf_padding=True,
             diffusion_steps=1000, noise_schedule="squaredcos_cap_v2", timestep_respacing="60",
             use_scale_shift_norm=True, resblock_updown=True, use_fp16=True, cache_text_emb=False, inpaint=False,
             super_res=False, chinese=True, sketch=False, class_balanced=False, sketch_classes=0,
             dtype=torch.bfloat16)
    o.update(over)
    return o


def model_and_diffusion_upsample(**over):
    o = model_and_diffusion_defaults(image_size=256, num_res_blocks=2, channel_mult=(1, 1, 2, 2, 4, 4),
                                     attention_resolutions=(32, 16, 8), dropout=0.0, n_vocab=50257,
                                     noise_schedule="linear", timestep_respacing="fast27", chinese=False)
    o.update(over)
    return o


# -------------------------------------------------------------------------------------------------------------
# schedules (host, float64 like the reference)
# -------------------------------------------------------------------------------------------------------------
def get_named_beta_schedule(schedule_name, num_diffusion_timesteps):
    """gaussian_computation.py:24-45.  "linear": Ho et al. betas rescaled to the step count; "squaredcos_cap_v2":
    beta_i = min(1 - abar((i+1)/T) / abar(i/T), 0.999) with abar(t) = cos^2((t + 0.008)/1.008 * pi/2)."""
    T = int(num_diffusion_timesteps)
    if schedule_name == "linear":
        k = 1000.0 / T
        return np.linspace(k * 1e-4, k * 0.02, T, dtype=np.float64)
    if schedule_name == "squaredcos_cap_v2":
        abar = lambda t: math.cos((t + 0.008) / 1.008 * math.pi / 2) ** 2
        return np.array([min(1.0 - abar((i + 1) / T) / abar(i / T), 0.999) for i in range(T)], dtype=np.float64)
    raise NotImplementedError("unknown beta schedule: %s" % schedule_name)


def space_timesteps(num_timesteps, section_counts):
    """gaussian_computation.py:103-160: the retained steps of the base process.  "ddimN": the integer stride that
    yields exactly N steps; "fast27": "10,10,3,2,2" with T-1 replaced by T-3; otherwise comma-separated counts, one
    per equal section, each section strided evenly (first and last step of the section kept)."""
    T = int(num_timesteps)
    if isinstance(section_counts, str):
        if section_counts.startswith("ddim"):
            want = int(section_counts[4:])
            for stride in range(1, T):
                if len(range(0, T, stride)) == want:
                    return set(range(0, T, stride))
            raise ValueError("cannot create exactly %d steps with an integer stride" % T)
        if section_counts == "fast27":
            steps = space_timesteps(T, "10,10,3,2,2")
            steps.remove(T - 1)
            steps.add(T - 3)
            return steps
        section_counts = [int(s) for s in section_counts.split(",")]
    base, extra = divmod(T, len(section_counts))
    kept, start = [], 0
    for i, count in enumerate(section_counts):
        size = base + (1 if i < extra else 0)
        if size < count:
            raise ValueError("cannot divide section of %d steps into %d" % (size, count))
        stride = 1 if count <= 1 else (size - 1) / (count - 1)
        pos = 0.0
        for _ in range(count):
            kept.append(start + round(pos))
            pos += stride
        start += size
    return set(kept)


def alpha_calculator(betas):
    betas = np.asarray(betas, dtype=np.float64)
    if betas.ndim != 1 or not ((betas > 0).all() and (betas <= 1).all()):
        raise ValueError("betas must be 1-D with values in (0, 1]")
    return np.cumprod(1.0 - betas)


def space_diffusion_from_base(use_timesteps, alphas_cumprod):
    """diffusion_creator.py:96-106: betas of the respaced process and the map back to base timesteps."""
    tmap, betas, last = [], [], 1.0
    for i, a in enumerate(alphas_cumprod):
        if i in use_timesteps:
            betas.append(1.0 - a / last)
            last = a
            tmap.append(i)
    return tmap, np.array(betas)


class DiffusionTables:
    """The per-timestep constants PMeanVariance.__init__ precomputes (gaussian_diffusion.py:133-157), including its
    float32 cumulative product (`self.betas = np.array(betas, dtype=np.float32)`, :133)."""

    def __init__(self, options):
        base = get_named_beta_schedule(options["noise_schedule"], options["diffusion_steps"])
        resp = options["timestep_respacing"] or [options["diffusion_steps"]]
        use = space_timesteps(options["diffusion_steps"], resp)
        self.timestep_map, new_betas = space_diffusion_from_base(use, alpha_calculator(base))
        b = np.array(new_betas, dtype=np.float32)
        self.betas = b
        self.num_timesteps = int(b.shape[0])
        ac = np.cumprod(1.0 - b, axis=0)
        ac_prev = np.append(1.0, ac[:-1])
        self.alphas_cumprod, self.alphas_cumprod_prev = ac, ac_prev
        self.sqrt_recip = np.sqrt(1.0 / ac)
        self.sqrt_recipm1 = np.sqrt(1.0 / ac - 1)
        pv = b * (1.0 - ac_prev) / (1.0 - ac)
        self.posterior_log_variance_clipped = np.log(np.append(pv[1], pv[1:]))
        self.coef1 = b * np.sqrt(ac_prev) / (1.0 - ac)
        self.coef2 = (1.0 - ac_prev) * np.sqrt(1.0 - b) / (1.0 - ac)
        self.log_betas = np.log(b)

    def step_coef(self, i, eta=0.0):
        """The 10 floats mdx_sampler_glide_step takes (include/mdx.h)."""
        f = lambda v: float(np.float32(v))
        return [f(self.sqrt_recip[i]), f(self.sqrt_recipm1[i]), f(self.coef1[i]), f(self.coef2[i]),
                f(self.posterior_log_variance_clipped[i]), f(self.log_betas[i]), f(self.alphas_cumprod[i]),
                f(self.alphas_cumprod_prev[i]), float(eta), 0.0 if i == 0 else 1.0]


# -------------------------------------------------------------------------------------------------------------
# structure and parameter inventory
# -------------------------------------------------------------------------------------------------------------
UNET_KEYS = ("num_channels", "num_res_blocks", "channel_mult", "num_heads", "num_head_channels", "num_heads_upsample",
             "attention_resolutions", "use_scale_shift_norm", "resblock_updown")


def glide_structure(o, in_channels):
    """Block list of model/unet.py:371-497.  Layers: ("conv", cin, cout) | ("res", cin, cout, None|"down"|"up") |
    ("attn", ch, heads)."""
    if not (o["use_scale_shift_norm"] and o["resblock_updown"]):
        raise ValueError("only use_scale_shift_norm=True, resblock_updown=True (both shipped option sets) are built")
    mc, nrb = o["num_channels"], o["num_res_blocks"]
    mults, attn_ds = list(o["channel_mult"]), set(o["attention_resolutions"])
    nhc, nh = o["num_head_channels"], o["num_heads"]
    nh_up = nh if o["num_heads_upsample"] == -1 else o["num_heads_upsample"]

    def attn(ch, heads):
        if nhc != -1:
            if ch % nhc:
                raise ValueError("q,k,v channels %d is not divisible by num_head_channels %d" % (ch, nhc))
            heads = ch // nhc
        return ("attn", ch, heads)

    ch = int(mults[0] * mc)
    inp, chans, ds = [[("conv", in_channels, ch)]], [ch], 1
    for level, mult in enumerate(mults):
        for _ in range(nrb):
            block = [("res", ch, int(mult * mc), None)]
            ch = int(mult * mc)
            if ds in attn_ds:
                block.append(attn(ch, nh))
            inp.append(block)
            chans.append(ch)
        if level != len(mults) - 1:
            inp.append([("res", ch, ch, "down")])
            chans.append(ch)
            ds *= 2
    mid = [("res", ch, ch, None), attn(ch, nh), ("res", ch, ch, None)]
    out = []
    for level, mult in list(enumerate(mults))[::-1]:
        for i in range(nrb + 1):
            ich = chans.pop()
            block = [("res", ch + ich, int(mc * mult), None)]
            ch = int(mc * mult)
            if ds in attn_ds:
                block.append(attn(ch, nh_up))
            if level and i == nrb:
                block.append(("res", ch, ch, "up"))
                ds //= 2
            out.append(block)
    return inp, mid, out


def _walk(inp, mid, out):
    for bi, block in enumerate(inp):
        for li, layer in enumerate(block):
            yield "input_blocks.%d.%d." % (bi, li), layer
    for li, layer in enumerate(mid):
        yield "middle_block.%d." % li, layer
    for bi, block in enumerate(out):
        for li, layer in enumerate(block):
            yield "output_blocks.%d.%d." % (bi, li), layer


def glide_param_shapes(o, in_channels, out_channels=6):
    """name -> shape, the names the reference's cells register (what load_param_into_net consumes)."""
    mc, W = o["num_channels"], o["xf_width"]
    ted = 4 * mc
    S = {"time_embed.0.weight": (ted, mc), "time_embed.0.bias": (ted,),
         "time_embed.2.weight": (ted, ted), "time_embed.2.bias": (ted,)}
    inp, mid, out = glide_structure(o, in_channels)
    for pre, layer in _walk(inp, mid, out):
        if layer[0] == "conv":
            S[pre + "conv.weight"] = (layer[2], layer[1], 3, 3)
            S[pre + "conv.bias"] = (layer[2],)
        elif layer[0] == "res":
            cin, cout = layer[1], layer[2]
            S[pre + "in_layers_0.gamma"] = (cin,)
            S[pre + "in_layers_0.beta"] = (cin,)
            S[pre + "in_layers_2.conv.weight"] = (cout, cin, 3, 3)
            S[pre + "in_layers_2.conv.bias"] = (cout,)
            S[pre + "emb_layers.1.weight"] = (2 * cout, ted)
            S[pre + "emb_layers.1.bias"] = (2 * cout,)
            S[pre + "out_layers_0.gamma"] = (cout,)
            S[pre + "out_layers_0.beta"] = (cout,)
            S[pre + "out_layers_3.conv.weight"] = (cout, cout, 3, 3)
            S[pre + "out_layers_3.conv.bias"] = (cout,)
            if cin != cout:
                S[pre + "skip_connection.conv.weight"] = (cout, cin, 1, 1)
                S[pre + "skip_connection.conv.bias"] = (cout,)
        else:
            ch = layer[1]
            S[pre + "norm.gamma"] = (ch,)
            S[pre + "norm.beta"] = (ch,)
            S[pre + "qkv.conv.weight"] = (3 * ch, ch, 1, 1)       # nn.Conv1d keeps a 4-D weight [out, in, 1, k]
            S[pre + "qkv.conv.bias"] = (3 * ch,)
            S[pre + "encoder_kv.conv.weight"] = (2 * ch, W, 1, 1)
            S[pre + "encoder_kv.conv.bias"] = (2 * ch,)
            S[pre + "proj_out.conv.weight"] = (ch, ch, 1, 1)
            S[pre + "proj_out.conv.bias"] = (ch,)
    c0 = int(o["channel_mult"][0] * mc)
    S["out.0.gamma"] = (c0,)
    S["out.0.beta"] = (c0,)
    S["out2.conv.weight"] = (out_channels, c0, 3, 3)
    S["out2.conv.bias"] = (out_channels,)
    for i in range(o["xf_layers"]):
        b = "transformer.resblocks.%d." % i
        S[b + "attn.c_qkv.weight"] = (3 * W, W)
        S[b + "attn.c_qkv.bias"] = (3 * W,)
        S[b + "attn.c_proj.weight"] = (W, W)
        S[b + "attn.c_proj.bias"] = (W,)
        S[b + "ln_1.gamma"] = (W,)
        S[b + "ln_1.beta"] = (W,)
        S[b + "mlp.c_fc.weight"] = (4 * W, W)
        S[b + "mlp.c_fc.bias"] = (4 * W,)
        S[b + "mlp.c_proj.weight"] = (W, 4 * W)
        S[b + "mlp.c_proj.bias"] = (W,)
        S[b + "ln_2.gamma"] = (W,)
        S[b + "ln_2.beta"] = (W,)
    if o["xf_final_ln"]:
        S["final_ln.gamma"] = (W,)
        S["final_ln.beta"] = (W,)
    S["token_embedding.embedding_table"] = (o["n_vocab"], W)
    S["positional_embedding"] = (o["text_ctx"], W)
    S["transformer_proj.weight"] = (ted, W)
    S["transformer_proj.bias"] = (ted,)
    if o["xf_padding"]:
        S["padding_embedding"] = (o["text_ctx"], W)
    return S


def random_glide_params(o, in_channels, seed=0, out_channels=6):
    """Seeded stand-in weights (no checkpoint can be downloaded here): fan-in scaled normals, norm gains near one,
    and NON-zero values for the reference's zero_module layers so that every path contributes to the output."""
    rng = np.random.default_rng(seed)
    P = {}
    for name, shape in glide_param_shapes(o, in_channels, out_channels).items():
        if name.endswith(".gamma"):
            v = 1.0 + 0.1 * rng.standard_normal(shape)
        elif name.endswith(".beta") or name.endswith(".bias"):
            v = 0.05 * rng.standard_normal(shape)
        elif name.endswith("embedding_table") or name.endswith("_embedding"):
            v = 0.5 * rng.standard_normal(shape)
        else:
            fan_in = int(np.prod(shape[1:]))
            gain = 0.5 if (".out_layers_3." in name or ".proj_out." in name or ".c_proj." in name) else 1.0
            v = gain * rng.standard_normal(shape) / math.sqrt(fan_in)
        P[name] = v.astype(np.float32)
    return P


# -------------------------------------------------------------------------------------------------------------
# the network
# -------------------------------------------------------------------------------------------------------------
def _head_perm(heads, ch, parts):
    """Row permutation from the reference's per-head layout [h][part][i] (QKVAttention's reshape + Split,
    unet.py:259-263; xf.py:86-88) to this build's fused layout [part][h][i]."""
    idx = np.empty(parts * heads * ch, dtype=np.int64)
    ar = np.arange(ch)
    for j in range(parts):
        for h in range(heads):
            idx[(j * heads + h) * ch:(j * heads + h + 1) * ch] = (h * parts + j) * ch + ar
    return torch.from_numpy(idx)


class Text2ImUNet:
    """model/text2im_model.py:24-121.  `model(x, timesteps, tokens, mask)` is the reference call; the sampler uses
    set_prompt / set_random_prompt / forward_nhwc so that per-prompt work is not repeated every step."""
    low_res_channels = 0

    def __init__(self, params, options, dtype=torch.bfloat16, device="cuda", prefix="", out_channels=6):
        if not torch.cuda.is_available():
            raise RuntimeError("minddiffusion_b200 needs a CUDA device (sm_100a); there is no CPU path")
        if dtype not in (torch.bfloat16, torch.float32):
            raise ValueError("dtype must be torch.bfloat16 (product) or torch.float32 (check mode)")
        self.o = dict(options)
        self.dtype, self.bf16 = dtype, dtype == torch.bfloat16
        self.device = torch.device(device)
        self.prefix = prefix
        self.in_channels = 3 + self.low_res_channels
        self.out_channels = out_channels
        self.model_channels = self.o["num_channels"]
        self.image_size = self.o["image_size"]
        self.text_ctx, self.xf_width = self.o["text_ctx"], self.o["xf_width"]
        self.inp, self.mid, self.out = glide_structure(self.o, self.in_channels)
        self.cin_pad = 64 if self.bf16 else self.in_channels
        self.cout_pad = 16 if self.bf16 else out_channels
        self.L = {}
        self._n = 0                 # images per forward the text buffers are sized for
        self._low_up = None
        self._graphs = {}
        self.use_cuda_graph = False
        self.replayed_launches = 0
        self._pack(params)

    # ---------------------------------------------------------------------------------------------------------
    def _w(self, w2d):
        return w2d.to(self.dtype).contiguous()

    def _pack(self, P):
        def g(name):
            key = self.prefix + name
            if key not in P:
                raise KeyError("missing parameter %s" % key)
            return _t(P[key], self.device)

        conv_k = lambda w: w.permute(0, 2, 3, 1).reshape(w.shape[0], -1)
        L, W = self.L, self.xf_width
        L["te0"] = (g("time_embed.0.weight"), g("time_embed.0.bias"))
        L["te2"] = (g("time_embed.2.weight"), g("time_embed.2.bias"))
        emb_w, emb_b, off = [], [], 0
        ekv_w, ekv_b, koff = [], [], 0
        ds = 1
        self.attn_layers = []
        for pre, layer in _walk(self.inp, self.mid, self.out):
            if layer[0] == "conv":
                w = g(pre + "conv.weight")
                if self.cin_pad != w.shape[1]:
                    wp = torch.zeros((w.shape[0], self.cin_pad, 3, 3), device=self.device)
                    wp[:, :w.shape[1]] = w
                    w = wp
                L[pre + "conv"] = (self._w(conv_k(w)), g(pre + "conv.bias"))
            elif layer[0] == "res":
                cin, cout, ud = layer[1], layer[2], layer[3]
                L[pre + "n1"] = (g(pre + "in_layers_0.gamma"), g(pre + "in_layers_0.beta"))
                if ud == "up" and self.bf16:  # nearest x2 + conv3x3 as four 2x2 phase GEMMs over the source image
                    L[pre + "c1"] = ([self._w(wp) for wp in fold_upsample_weights(g(pre + "in_layers_2.conv.weight"))],
                                     g(pre + "in_layers_2.conv.bias"))
                else:
                    L[pre + "c1"] = (self._w(conv_k(g(pre + "in_layers_2.conv.weight"))),
                                     g(pre + "in_layers_2.conv.bias"))
                emb_w.append(g(pre + "emb_layers.1.weight"))
                emb_b.append(g(pre + "emb_layers.1.bias"))
                L[pre + "emb"] = off
                off += 2 * cout
                L[pre + "n2"] = (g(pre + "out_layers_0.gamma"), g(pre + "out_layers_0.beta"))
                w2, b2 = conv_k(g(pre + "out_layers_3.conv.weight")), g(pre + "out_layers_3.conv.bias")
                if cin != cout:
                    ws = g(pre + "skip_connection.conv.weight").reshape(cout, cin)
                    bs = g(pre + "skip_connection.conv.bias")
                    if self.bf16:  # the 1x1 skip conv rides behind the 3x3 taps in the same accumulator
                        L[pre + "c2"] = (self._w(torch.cat((w2, ws), dim=1)), (b2 + bs).contiguous(), True)
                    else:
                        L[pre + "c2"] = (self._w(w2), b2, True)
                        L[pre + "skip"] = (self._w(ws), bs)
                else:
                    L[pre + "c2"] = (self._w(w2), b2, False)
                if ud == "down":
                    ds *= 2
                elif ud == "up":
                    ds //= 2
            else:
                ch, heads = layer[1], layer[2]
                d = ch // heads
                L[pre + "norm"] = (g(pre + "norm.gamma"), g(pre + "norm.beta"))
                p3 = _head_perm(heads, d, 3).to(self.device)
                L[pre + "qkv"] = (self._w(g(pre + "qkv.conv.weight").reshape(3 * ch, ch)[p3]),
                                  g(pre + "qkv.conv.bias")[p3].contiguous())
                p2 = _head_perm(heads, d, 2).to(self.device)
                ekv_w.append(g(pre + "encoder_kv.conv.weight").reshape(2 * ch, W)[p2])
                ekv_b.append(g(pre + "encoder_kv.conv.bias")[p2])
                L[pre + "proj"] = (self._w(g(pre + "proj_out.conv.weight").reshape(ch, ch)), g(pre + "proj_out.conv.bias"))
                res = self.image_size // ds
                self.attn_layers.append((pre, ch, koff, res * res))
                koff += 2 * ch
        L["emb_all"] = (torch.cat(emb_w, dim=0).contiguous(), torch.cat(emb_b, dim=0).contiguous())
        L["ekv_all"] = (self._w(torch.cat(ekv_w, dim=0)), torch.cat(ekv_b, dim=0).contiguous())
        L["out_norm"] = (g("out.0.gamma"), g("out.0.beta"))
        w, b = conv_k(g("out2.conv.weight")), g("out2.conv.bias")
        if self.cout_pad != w.shape[0]:
            wp = torch.zeros((self.cout_pad, w.shape[1]), device=self.device)
            wp[:w.shape[0]] = w
            bp = torch.zeros(self.cout_pad, device=self.device)
            bp[:b.shape[0]] = b
            w, b = wp, bp
        L["out_conv"] = (self._w(w), b.contiguous())
        # text transformer (xf.py) -- c_qkv rows re-ordered to [Q | K | V]
        heads = self.o["xf_heads"]
        pq = _head_perm(heads, W // heads, 3).to(self.device)
        self.xf = []
        for i in range(self.o["xf_layers"]):
            b_ = "transformer.resblocks.%d." % i
            self.xf.append(dict(
                ln1=(g(b_ + "ln_1.gamma"), g(b_ + "ln_1.beta")), ln2=(g(b_ + "ln_2.gamma"), g(b_ + "ln_2.beta")),
                qkv=(self._w(g(b_ + "attn.c_qkv.weight")[pq]), g(b_ + "attn.c_qkv.bias")[pq].contiguous()),
                out=(self._w(g(b_ + "attn.c_proj.weight")), g(b_ + "attn.c_proj.bias")),
                fc=(self._w(g(b_ + "mlp.c_fc.weight")), g(b_ + "mlp.c_fc.bias")),
                proj=(self._w(g(b_ + "mlp.c_proj.weight")), g(b_ + "mlp.c_proj.bias"))))
        self.final_ln = (g("final_ln.gamma"), g("final_ln.beta")) if self.o["xf_final_ln"] else None
        self.table, self.pos = g("token_embedding.embedding_table"), g("positional_embedding")
        self.pad = g("padding_embedding") if self.o["xf_padding"] else None
        L["tproj"] = (self._w(g("transformer_proj.weight")), g("transformer_proj.bias"))

    # ---------------------------------------------------------------------------------------------------------
    def _dense(self, x, w, b, residual=None, act=None, trunk=False):
        if self.bf16:
            return ops.gemm_tc([(x, 1)], w, bias=b, residual=residual, act=act,
                               out_dtype=torch.float32 if trunk else torch.bfloat16)
        return ops.conv_direct(x, w, bias=b, residual=residual, act=act)

    def _conv3(self, x, w, b, residual=None, skips=None, trunk=False):
        n, H, W_, _ = x.shape
        if self.bf16:
            segs = [(x, 9)] + [(s, 1) for s in (skips or [])]
            return ops.gemm_tc(segs, w, bias=b, residual=residual,
                               out_dtype=torch.float32 if trunk else torch.bfloat16).view(n, H, W_, -1)
        return ops.conv_direct(x, w, 3, 3, 1, 1, bias=b, residual=residual)

    # ---------------------------------------------------------------------------------------------------------
    # text side
    # ---------------------------------------------------------------------------------------------------------
    def encode_text(self, tokens, mask):
        """get_text_emb (text2im_model.py:79-91).  tokens/mask: int [B, L].  Returns (xf_proj fp32 [B, 4*mc],
        xf_out [B*L, width] in the operand dtype, row-major tokens: the reference's NCL transpose is a layout detail
        of its Conv1d encoder_kv)."""
        dt, W, heads = self.dtype, self.xf_width, self.o["xf_heads"]
        ids = tokens.to(device=self.device, dtype=torch.int32).contiguous()
        B, Lt = ids.shape
        if Lt != self.text_ctx:
            raise ValueError("expected %d tokens per prompt, got %d" % (self.text_ctx, Lt))
        m = None
        if self.pad is not None:
            if mask is None:
                raise ValueError("xf_padding=True needs the token mask")
            m = mask.to(device=self.device, dtype=torch.int32).contiguous()
        x = ops.embedding_masked(self.table, self.pos, self.pad, ids, m)          # fp32 token stream [B*L, W]
        scale = (W // heads) ** -0.5                                              # xf.py:80: d^-1/4 on q and on k
        for blk in self.xf:
            h = ops.layernorm(x, *blk["ln1"], eps=1e-7, out_dtype=dt)             # nn.LayerNorm default epsilon
            qkv = self._dense(h, *blk["qkv"]).view(B, Lt, 3 * W)
            o = ops.attention(qkv[:, :, :W], qkv[:, :, W:2 * W], qkv[:, :, 2 * W:], heads, scale)
            x = self._dense(o.view(B * Lt, W), *blk["out"], residual=x, trunk=True)
            h = ops.layernorm(x, *blk["ln2"], eps=1e-7, out_dtype=dt)
            h = self._dense(h, *blk["fc"], act="gelu")
            x = self._dense(h, *blk["proj"], residual=x, trunk=True)
        if self.final_ln is not None:
            xf_out = ops.layernorm(x, *self.final_ln, eps=1e-7, out_dtype=dt)
        else:
            xf_out = ops.cast(x, dt) if self.bf16 else x
        last = xf_out.view(B, Lt, W)[:, -1]                                        # strided [B, W] view
        xf_proj = self._dense(last, *self.L["tproj"], trunk=True)
        return xf_proj, xf_out

    def _alloc_text(self, n):
        if self._n == n:
            return
        self._n = n
        self._graphs = {}
        Lt = self.text_ctx
        self._xfp = torch.zeros((n, 4 * self.model_channels), dtype=torch.float32, device=self.device)
        self._kv = {pre: torch.zeros((n, Lt + N, 2 * ch), dtype=self.dtype, device=self.device)
                    for pre, ch, _, N in self.attn_layers}

    def _fill_text(self, lo, hi, xf_proj, xf_out):
        """Write the text-dependent operands of images [lo, hi): the transformer_proj rows added to the time
        embedding and every AttentionBlock's encoder K/V rows (unet.py:236-238).  One text row broadcasts."""
        B = xf_proj.shape[0]
        Lt = self.text_ctx
        ops.copy_rows(xf_proj.view(B, 1, -1), self._xfp[lo:hi].view(hi - lo, 1, -1))
        ekv = self._dense(xf_out, *self.L["ekv_all"]).view(B, Lt, -1)
        for pre, ch, koff, _ in self.attn_layers:
            ops.copy_rows(ekv[:, :, koff:koff + 2 * ch], self._kv[pre][lo:hi, :Lt, :])

    def set_prompt(self, tokens, mask, n_total=None):
        """Text operands of images [0, B) for B prompts; n_total = images per forward (2B under guidance)."""
        B = tokens.shape[0]
        self._alloc_text(n_total or B)
        self._fill_text(0, B, *self.encode_text(tokens, mask))

    def set_random_prompt(self, token_row, mask_row, lo, hi):
        """CombinePrompt's second half (guider.py:36-41): ONE prompt row shared by images [lo, hi)."""
        xf_proj, xf_out = self.encode_text(token_row.view(1, -1), None if mask_row is None else mask_row.view(1, -1))
        self._fill_text(lo, hi, xf_proj, xf_out)

    # ---------------------------------------------------------------------------------------------------------
    # image side
    # ---------------------------------------------------------------------------------------------------------
    def _res(self, pre, layer, x1, x2, emb_all):
        """ResBlock.construct (unet.py:166-201) with use_scale_shift_norm; x1 (and x2, the skip to concatenate) are
        fp32 residual-stream tensors in bf16 mode."""
        L, dt = self.L, self.dtype
        cout, ud = layer[2], layer[3]
        g1, b1 = L[pre + "n1"]
        w2, bb2, has_skip = L[pre + "c2"]
        raw = None
        if self.bf16 and has_skip:
            h, raw = ops.groupnorm(x1, g1, b1, 32, 1e-5, silu=True, x2=x2, scratch=self._gn_scratch, out_dtype=dt,
                                   raw_copy=True)
        else:
            h = ops.groupnorm(x1, g1, b1, 32, 1e-5, silu=True, x2=x2, scratch=self._gn_scratch, out_dtype=dt)
        x = x1
        if ud == "down":       # h_upd / x_upd = AvgPool2d(2, 2) (unet.py:139-141)
            h, x = ops.avgpool2x2(h), ops.avgpool2x2(x1)
        elif ud == "up":       # nearest x2, no conv (unet.py:49-52, 136-138)
            x = ops.upsample_nearest2x(x1)
            if not self.bf16:
                h = ops.upsample_nearest2x(h)
        if ud == "up" and self.bf16:
            h = upsample_conv3x3(h, L[pre + "c1"][0], L[pre + "c1"][1], self.dtype)
        else:
            h = self._conv3(h, *L[pre + "c1"])
        off = L[pre + "emb"]
        g2, b2 = L[pre + "n2"]
        h = ops.groupnorm_film(h, g2, b2, emb_all[:, off:off + 2 * cout], 32, 1e-5, silu=True,
                               scratch=self._gn_scratch)
        if not has_skip:
            return self._conv3(h, w2, bb2, residual=x, trunk=True)
        if self.bf16:
            return self._conv3(h, w2, bb2, skips=[raw], trunk=True)
        xin = x1 if x2 is None else ops.concat_channels(x1, x2)
        ws, bs = L[pre + "skip"]
        return self._conv3(h, w2, bb2, residual=ops.conv_direct(xin, ws, bias=bs))

    def _attn(self, pre, layer, x):
        """AttentionBlock.construct + QKVAttention (unet.py:231-277): keys/values are [encoder K/V | spatial K/V];
        the reference scales q and k by ch^-1/4 each, i.e. the logits by ch^-1/2."""
        L, dt = self.L, self.dtype
        ch, heads = layer[1], layer[2]
        n, H, W_, _ = x.shape
        N, Lt = H * W_, self.text_ctx
        kv = self._kv[pre]
        if kv.shape[1] != Lt + N:
            raise ValueError("%s: text buffers were sized for %d positions, got %d (image_size mismatch)"
                             % (pre, kv.shape[1] - Lt, N))
        hn = ops.groupnorm(x, *L[pre + "norm"], 32, 1e-5, silu=False, scratch=self._gn_scratch, out_dtype=dt)
        qkv = self._dense(hn.view(n * N, ch), *L[pre + "qkv"]).view(n, N, 3 * ch)
        ops.copy_rows(qkv[:, :, ch:], kv[:, Lt:, :])
        o = ops.attention(qkv[:, :, :ch], kv[:, :, :ch], kv[:, :, ch:], heads, (ch // heads) ** -0.5)
        return self._dense(o.view(n * N, ch), *L[pre + "proj"], residual=x.view(n * N, ch), trunk=True).view(n, H, W_, ch)

    def _forward_impl(self, x_nchw, t, repeat, taps=None):
        L = self.L
        n = x_nchw.shape[0] * repeat
        self._gn_scratch = torch.empty(n * 64, dtype=torch.float64, device=self.device)
        t_emb = ops.timestep_embedding(t, self.model_channels)                 # simple_nn.py:158-164: [cos | sin]
        emb = ops.conv_direct(t_emb, L["te0"][0], bias=L["te0"][1])
        emb = ops.conv_direct(emb, L["te2"][0], bias=L["te2"][1], pre_silu=True, residual=self._xfp)  # + xf_proj
        emb_all = ops.conv_direct(emb, L["emb_all"][0], bias=L["emb_all"][1], pre_silu=True)
        h = ops.nchw2_to_nhwc(x_nchw, self._low_up, self.dtype, repeat=repeat, cpad=self.cin_pad)
        hs = []

        def run(pre, layer, h, skip):
            if layer[0] == "conv":
                return self._conv3(h, *L[pre + "conv"], trunk=True)
            if layer[0] == "res":
                return self._res(pre, layer, h, skip, emb_all)
            return self._attn(pre, layer, h)

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
        h = ops.groupnorm(h, *L["out_norm"], 32, 1e-5, silu=True, scratch=self._gn_scratch, out_dtype=self.dtype)
        w, b = L["out_conv"]
        return self._conv3(h, w, b, trunk=True)

    def forward_nhwc(self, x_nchw, timesteps, repeat=1, taps=None, random_prompt=None):
        """x_nchw: fp32 [B, 3, H, W], read `repeat` times (2 = CombinePrompt's concat((half, half)), guider.py:32-35);
        timesteps: [B*repeat] mapped timesteps.  Text operands must have been set for B*repeat images.  Returns fp32
        NHWC [B*repeat, H, W, ld >= out_channels].
        random_prompt = (token_row, mask_row, lo, hi): run set_random_prompt for images [lo, hi) first — inside the
        same captured graph when graphs are on, so that a guided step is ONE graph launch."""
        n = x_nchw.shape[0] * repeat
        if self._n != n:
            raise RuntimeError("text operands are set for %d images, forward has %d (call set_prompt)" % (self._n, n))
        if self.low_res_channels and self._low_up is None:
            raise RuntimeError("call set_low_res(samples) before forward")
        x_nchw = x_nchw.to(device=self.device, dtype=torch.float32).contiguous()
        t = timesteps.to(device=self.device, dtype=torch.float32).contiguous()
        if t.numel() != n:
            raise ValueError("timesteps must have %d entries" % n)
        if random_prompt is not None:
            tok, msk, lo, hi = random_prompt
            tok = tok.to(device=self.device, dtype=torch.int32).contiguous()
            msk = (torch.ones_like(tok) if msk is None else msk.to(device=self.device, dtype=torch.int32)).contiguous()
        if not self.use_cuda_graph or taps is not None:
            if random_prompt is not None:
                self.set_random_prompt(tok, msk, lo, hi)
            return self._forward_impl(x_nchw, t, repeat, taps)
        key = (tuple(x_nchw.shape), repeat, None if random_prompt is None else (lo, hi))
        ent = self._graphs.get(key)
        if ent is None:
            sx, st_ = x_nchw.clone(), t.clone()
            stok, smsk = (tok.clone(), msk.clone()) if random_prompt is not None else (None, None)

            def body():
                if stok is not None:
                    self.set_random_prompt(stok, smsk, lo, hi)
                return self._forward_impl(sx, st_, repeat)

            body()
            torch.cuda.synchronize()
            graph = torch.cuda.CUDAGraph()
            n0 = _lib.launch_count()
            with torch.cuda.graph(graph):
                sout = body()
            ent = (graph, sx, st_, stok, smsk, sout, _lib.launch_count() - n0)
            self._graphs[key] = ent
        graph, sx, st_, stok, smsk, sout, nk = ent
        sx.copy_(x_nchw)
        st_.copy_(t)
        if stok is not None:
            stok.copy_(tok)
            smsk.copy_(msk)
        graph.replay()
        self.replayed_launches += nk
        return sout

    def __call__(self, x, timesteps, tokens=None, mask=None):
        """Reference call (text2im_model.py:93): NCHW in, NCHW [N, out_channels, H, W] out."""
        self.set_prompt(tokens, mask)
        out = self.forward_nhwc(x, timesteps)
        return ops.nhwc_to_nchw(out.contiguous(), self.out_channels)


class SuperResText2ImUNet(Text2ImUNet):
    """model/text2im_model.py:124-238: the UNet input is concat(x, bilinear upsample of the low-res sample)."""
    low_res_channels = 3

    def set_low_res(self, low_res, quantize_u8=False):
        """low_res: fp32 NCHW [B, 3, h, w] -> the conditioning channels, resized to image_size (text2im_model.py:191).
        quantize_u8 applies PMeanVariance.superResPreprocess first (the sampler's path)."""
        low = low_res.to(device=self.device, dtype=torch.float32).contiguous()
        up = ops.resize_bilinear(low, self.image_size, self.image_size, quantize_u8)
        if self._low_up is not None and self._low_up.shape == up.shape:
            self._low_up.copy_(up)   # same shape: refill the buffer the captured graphs read
        else:
            self._low_up = up
            self._graphs = {}

    def __call__(self, x, timesteps, low_res=None, tokens=None, mask=None):
        self.set_low_res(low_res)
        return super().__call__(x, timesteps, tokens, mask)


def create_model(params, device="cuda", **options):
    """model_creator.py:18-75 (options = model_and_diffusion_defaults(); options["dtype"] picks the mode)."""
    return Text2ImUNet(params, options, dtype=options.get("dtype", torch.bfloat16), device=device)


def create_upsample_model(params, device="cuda", **options):
    """model_creator.py:78-135 (options = model_and_diffusion_upsample())."""
    return SuperResText2ImUNet(params, options, dtype=options.get("dtype", torch.bfloat16), device=device)


# -------------------------------------------------------------------------------------------------------------
# samplers (model/gaussian_diffusion.py, model/guider.py) and loops (main_funcs.py)
# -------------------------------------------------------------------------------------------------------------
class _DiffusionBase:
    def __init__(self, model, options, shape):
        self.model = model
        self.tables = DiffusionTables(options)
        self.shape = tuple(shape)
        self.device = model.device
        self._prompt_key = None
        self.tmap = torch.tensor(self.tables.timestep_map, dtype=torch.float32, device=self.device)

    @property
    def num_timesteps(self):
        return self.tables.num_timesteps

    def _index(self, timesteps):
        i = int(timesteps.reshape(-1)[0]) if isinstance(timesteps, torch.Tensor) else int(timesteps)
        if not 0 <= i < self.num_timesteps:
            raise IndexError("timestep %d outside the respaced process (0..%d)" % (i, self.num_timesteps - 1))
        return i

    def _prompt_changed(self, token, mask):
        key = (id(token), token._version, id(mask), None if mask is None else mask._version)
        changed = key != self._prompt_key
        self._prompt_key = key
        return changed


class GenerativePSampleDiffusionModel(_DiffusionBase):
    """init_diffusion_model's product (diffusion_creator.py:26-39): SamplingWithGuidance(Text2ImUNet) ->
    PMeanVariance -> PSample.  shape = (2 * pics_generated, 3, S, S) as in the reference; only the first half of
    x is read (CombinePrompt, guider.py:32-35) and the returned sample repeats the first half, because the second
    half of the reference's return value is a by-product that txt2img.py:121 slices away."""

    def __init__(self, model, options, guidance_scale, shape):
        super().__init__(model, options, shape)
        self.guidance_scale = float(guidance_scale)
        self.P = self.shape[0] // 2

    def set_prompt(self, token, mask):
        P = self.P
        self.model.set_prompt(token[:P], None if mask is None else mask[:P], n_total=2 * P)

    def step(self, x, i, random_token, random_mask=None, noise=None):
        """One guided ancestral step on x [P, 3, S, S]; returns (x_prev, pred_xstart)."""
        P, m = self.P, self.model
        t = self.tmap[i].expand(2 * P)
        out = m.forward_nhwc(x, t, repeat=2, random_prompt=(random_token, random_mask, P, 2 * P))
        if noise is None and i != 0:
            noise = torch.randn(x.shape, dtype=torch.float32, device=self.device)
        x0 = torch.empty_like(x)
        xp = ops.sampler_glide_step(out, x, self.tables.step_coef(i), True, self.guidance_scale, 0, noise=noise,
                                    pred_xstart=x0)
        return xp, x0

    def __call__(self, x, timesteps, token, mask, random_token=None, random_mask=None, is_train=False, noise=None):
        if self._prompt_changed(token, mask):
            self.set_prompt(token, mask)
        P = self.P
        xh = x[:P].to(device=self.device, dtype=torch.float32).contiguous()
        xp, x0 = self.step(xh, self._index(timesteps), random_token.to(self.device),
                           None if random_mask is None else random_mask.to(self.device),
                           None if noise is None else noise[:P].to(self.device, torch.float32).contiguous())
        return torch.cat((xp, xp)), torch.cat((x0, x0))


class DDimSampleDiffusionModel(_DiffusionBase):
    """init_super_res_model's product (diffusion_creator.py:42-56): SuperResText2ImUNet -> PMeanVariance ->
    DDimSample (eta = 0).  No guidance: every image carries its own prompt."""

    def __init__(self, model, options, shape, eta=0.0):
        super().__init__(model, options, shape)
        self.eta = float(eta)
        self._samples_key = None

    def set_prompt(self, token, mask, samples):
        self.model.set_prompt(token, mask)
        self.model.set_low_res(samples, quantize_u8=True)

    def step(self, x, i, noise=None):
        m = self.model
        out = m.forward_nhwc(x, self.tmap[i].expand(x.shape[0]))
        if noise is None and i != 0 and self.eta != 0.0:
            noise = torch.randn(x.shape, dtype=torch.float32, device=self.device)
        x0 = torch.empty_like(x)
        xp = ops.sampler_glide_step(out, x, self.tables.step_coef(i, self.eta), False, 1.0, 1, noise=noise,
                                    pred_xstart=x0)
        return xp, x0

    def __call__(self, x, timesteps, token, mask, samples, is_train=False, noise=None):
        skey = (id(samples), samples._version)
        if self._prompt_changed(token, mask) or skey != self._samples_key:
            self._samples_key = skey
            self.set_prompt(token, mask, samples)
        xh = x.to(device=self.device, dtype=torch.float32).contiguous()
        return self.step(xh, self._index(timesteps), noise)


def init_diffusion_model(options, guidance_scale, shape, params=None, ckpt_path=None, device="cuda"):
    """diffusion_creator.py:26-39.  params: name -> array (the reference loads a MindSpore checkpoint afterwards,
    txt2img.py:101-102; here the weights are handed in, random stand-ins when None)."""
    if params is None:
        params = random_glide_params(options, 3)
    model = create_model(params, device=device, **options)
    return GenerativePSampleDiffusionModel(model, options, guidance_scale, shape)


def init_super_res_model(options, shape, params=None, ckpt_path=None, device="cuda"):
    """diffusion_creator.py:42-56."""
    if params is None:
        params = random_glide_params(options, 6)
    model = create_upsample_model(params, device=device, **options)
    return DDimSampleDiffusionModel(model, options, shape)


def gaussian_p_sample_loop(diffusion_model, token, mask, shape, num_timesteps, tokenizer=None, text_ctx=128,
                           noise=None, progress=False, dtype=None, vocab_len=50001, random_tokens=None,
                           step_noise=None):
    """main_funcs.py:20-41.  Returns fp32 [shape[0], 3, S, S] (first half = the images; see the class docstring).
    random_tokens [num_timesteps, text_ctx] / step_noise [num_timesteps, P, 3, S, S], indexed by timestep, replace
    the per-step draws (main_funcs.py:33, gaussian_diffusion.py:53) — a hook for parity tests."""
    dm, dev = diffusion_model, diffusion_model.device
    P = dm.P
    img = noise if noise is not None else torch.randn(tuple(shape), dtype=torch.float32, device=dev)
    x = img[:P].to(device=dev, dtype=torch.float32).contiguous()
    dm.set_prompt(token.to(dev), None if mask is None else mask.to(dev))
    ones = torch.ones((text_ctx,), dtype=torch.int32, device=dev)
    for i in range(num_timesteps - 1, -1, -1):
        if random_tokens is not None:
            rt = random_tokens[i].to(device=dev, dtype=torch.int32)
        else:
            rt = torch.randint(1, vocab_len - 1, (text_ctx,), dtype=torch.int32, device=dev)
        nz = None if step_noise is None else step_noise[i].to(device=dev, dtype=torch.float32).contiguous()
        x, _ = dm.step(x, i, rt, ones, nz)
    return torch.cat((x, x))


def ddim_sample_loop(super_res_model, up_shape, samples, token, mask, num_timesteps, noise=None, progress=False,
                     dtype=None):
    """main_funcs.py:44-65: x_T = 0.997 * N(0, I) unless given."""
    dm, dev = super_res_model, super_res_model.device
    if noise is not None:
        x = noise.to(device=dev, dtype=torch.float32).contiguous()
    else:
        x = torch.randn(tuple(up_shape), dtype=torch.float32, device=dev) * 0.997
    dm.set_prompt(token.to(dev), None if mask is None else mask.to(dev), samples)
    for i in range(num_timesteps - 1, -1, -1):
        x, _ = dm.step(x, i)
    return x
