"""ORACLE (test infrastructure only — never imported by the product path).

CPU restatement, in plain torch fp32 / numpy, of Taichu-GLIDE's denoising path
(`/root/reference/vision/Taichu-GLIDE/model/glide_text2im/`): model/unet.py, model/text2im_model.py, model/xf.py,
model/simple_nn.py, model/guider.py, model/gaussian_diffusion.py, gaussian_computation.py, diffusion_creator.py,
main_funcs.py.  Each function cites the lines it follows.  Everything is NCHW / [N, C, T] like the reference,
including its per-head [q|k|v] channel interleave, so a layout mistake in the CUDA path cannot hide here.

MindSpore semantics encoded explicitly: `_Linear` is y = x W^T + b with W [out, in]; nn.Conv1d/2d(pad_mode='pad');
nn.GroupNorm eps 1e-5 / nn.LayerNorm default epsilon 1e-7, biased variance; nn.GELU() is the tanh approximation;
nn.Dropout is the identity at inference; ops.ResizeNearestNeighbor maps dst -> floor(dst/2);
ops.ResizeBilinear(align_corners=False) keeps its default half_pixel_centers=False (src = dst * in/out);
ops.Round rounds half to even.

Pinning status: the reference ships no golden vectors for this family; the restatement is pinned against the
UNMODIFIED reference classes executed on the torch-backed `mindspore` stand-in (tests/golden/ms_shim.py,
tests/golden/make_golden_glide.py -> tests/golden/glide_golden.npz), checked by tests/test_glide.py.  The bilinear
resize and Round conventions above are MindSpore-documented behaviour that the stand-in restates as well, so those
two are pinned to the documentation, not to an executed MindSpore.
"""
import math

import numpy as np
import torch
import torch.nn.functional as F


# ---------------------------------------------------------------------------------------------------------
# schedules
# ---------------------------------------------------------------------------------------------------------
def named_betas(name, T):
    """gaussian_computation.py:24-62."""
    if name == "linear":
        s = 1000 / T
        return np.linspace(s * 0.0001, s * 0.02, T, dtype=np.float64)
    if name == "squaredcos_cap_v2":
        f = lambda t: math.cos((t + 0.008) / 1.008 * math.pi / 2) ** 2
        return np.array([min(1 - f((i + 1) / T) / f(i / T), 0.999) for i in range(T)])
    raise NotImplementedError(name)


def spaced_steps(T, spec):
    """gaussian_computation.py:103-160."""
    if isinstance(spec, str):
        if spec.startswith("ddim"):
            n = int(spec[4:])
            for s in range(1, T):
                if len(range(0, T, s)) == n:
                    return set(range(0, T, s))
            raise ValueError("no integer stride")
        if spec == "fast27":
            st = spaced_steps(T, "10,10,3,2,2")
            st.remove(T - 1)
            st.add(T - 3)
            return st
        spec = [int(x) for x in spec.split(",")]
    per, extra = T // len(spec), T % len(spec)
    start, res = 0, []
    for i, cnt in enumerate(spec):
        size = per + (1 if i < extra else 0)
        if size < cnt:
            raise ValueError("section too small")
        stride = 1 if cnt <= 1 else (size - 1) / (cnt - 1)
        cur = 0.0
        for _ in range(cnt):
            res.append(start + round(cur))
            cur += stride
        start += size
    return set(res)


class Tables:
    """create_gaussian_diffusion (diffusion_creator.py:59-95) + PMeanVariance.__init__ (gaussian_diffusion.py:133-157)."""

    def __init__(self, noise_schedule, diffusion_steps, timestep_respacing):
        base = named_betas(noise_schedule, diffusion_steps)
        use = spaced_steps(diffusion_steps, timestep_respacing or [diffusion_steps])
        ac64 = np.cumprod(1.0 - np.array(base, dtype=np.float64), axis=0)
        self.timestep_map, nb, last = [], [], 1.0
        for i, a in enumerate(ac64):
            if i in use:
                nb.append(1 - a / last)
                last = a
                self.timestep_map.append(i)
        self.betas = np.array(np.array(nb), dtype=np.float32)
        self.T = len(self.betas)
        self.ac = np.cumprod(1.0 - self.betas, axis=0)
        self.ac_prev = np.append(1.0, self.ac[:-1])
        f32 = lambda a: torch.tensor(np.asarray(a), dtype=torch.float32)
        self.sqrt_recip = f32(np.sqrt(1.0 / self.ac))
        self.sqrt_recipm1 = f32(np.sqrt(1.0 / self.ac - 1))
        pv = self.betas * (1.0 - self.ac_prev) / (1.0 - self.ac)
        self.post_logvar = f32(np.log(np.append(pv[1], pv[1:])))
        self.coef1 = f32(self.betas * np.sqrt(self.ac_prev) / (1.0 - self.ac))
        self.coef2 = f32((1.0 - self.ac_prev) * np.sqrt(1.0 - self.betas) / (1.0 - self.ac))
        self.log_betas = torch.log(f32(self.betas))
        self.ac_t, self.ac_prev_t = f32(self.ac), f32(self.ac_prev)


# ---------------------------------------------------------------------------------------------------------
# network
# ---------------------------------------------------------------------------------------------------------
def timestep_embedding(t, dim):
    """simple_nn.py:158-164."""
    half = dim // 2
    freqs = torch.exp(-math.log(10000) * torch.arange(0, half, dtype=torch.float32) / half)
    args = t.to(torch.float32)[:, None] * freqs[None]
    return torch.cat((torch.cos(args), torch.sin(args)), dim=-1)


def gn32(P, pre, x, swish):
    """simple_nn.py:32-58 (fp32 GroupNorm, 32 groups, eps 1e-5, optional SiLU)."""
    y = F.group_norm(x.float(), 32, P[pre + "gamma"], P[pre + "beta"], 1e-5)
    return F.silu(y) if swish else y


def conv(P, pre, x, pad):
    w = P[pre + "conv.weight"]
    if x.dim() == 3:  # nn.Conv1d stores [out, in, 1, k]
        return F.conv1d(x, w.reshape(w.shape[0], w.shape[1], -1), P[pre + "conv.bias"], padding=pad)
    return F.conv2d(x, w, P[pre + "conv.bias"], padding=pad)


def linear(P, pre, x):
    return F.linear(x, P[pre + "weight"], P[pre + "bias"])


def res_block(P, pre, x, emb, cin, cout, updown):
    """unet.py:166-201 with use_scale_shift_norm=True."""
    h = gn32(P, pre + "in_layers_0.", x, True)
    if updown == "up":      # unet.py:49-52: nearest, no conv
        h = F.interpolate(h, scale_factor=2, mode="nearest")
        x = F.interpolate(x, scale_factor=2, mode="nearest")
    elif updown == "down":  # unet.py:73: AvgPool2d(2, 2)
        h = F.avg_pool2d(h, 2, 2)
        x = F.avg_pool2d(x, 2, 2)
    h = conv(P, pre + "in_layers_2.", h, 1)
    eo = linear(P, pre + "emb_layers.1.", F.silu(emb))[..., None, None]
    scale, shift = torch.chunk(eo, 2, dim=1)
    h = gn32(P, pre + "out_layers_0.", h, False) * (1 + scale) + shift
    h = conv(P, pre + "out_layers_3.", F.silu(h), 1)
    if cin == cout:
        return x + h
    return conv(P, pre + "skip_connection.", x, 0) + h


def qkv_attention(qkv, heads, encoder_kv=None):
    """unet.py:251-277."""
    bs, width, length = qkv.shape
    ch = width // (3 * heads)
    q, k, v = torch.chunk(qkv.reshape(bs * heads, ch * 3, length), 3, dim=1)
    if encoder_kv is not None:
        ek, ev = torch.chunk(encoder_kv.reshape(bs * heads, ch * 2, -1), 2, dim=1)
        k = torch.cat((ek, k), dim=-1)
        v = torch.cat((ev, v), dim=-1)
    scale = 1 / math.sqrt(math.sqrt(ch))
    w = torch.matmul((q * scale).transpose(1, 2), k * scale)
    w = torch.softmax(w.float(), dim=-1)
    a = torch.matmul(w, v.transpose(1, 2)).transpose(1, 2)
    return a.reshape(bs, -1, length)


def attention_block(P, pre, x, heads, encoder_out):
    """unet.py:231-241."""
    b, c, d, e = x.shape
    qkv = conv(P, pre + "qkv.", gn32(P, pre + "norm.", x, False).reshape(b, c, -1), 0)
    ekv = conv(P, pre + "encoder_kv.", encoder_out, 0)
    h = conv(P, pre + "proj_out.", qkv_attention(qkv, heads, ekv), 0)
    return x + h.reshape(b, c, d, e)


def structure(o, in_channels):
    """unet.py:371-497."""
    mc, nrb = o["num_channels"], o["num_res_blocks"]
    mults, attn_ds = list(o["channel_mult"]), set(o["attention_resolutions"])

    def heads(ch, nh):
        return nh if o["num_head_channels"] == -1 else ch // o["num_head_channels"]

    nh = o["num_heads"]
    nhu = nh if o["num_heads_upsample"] == -1 else o["num_heads_upsample"]
    ch = int(mults[0] * mc)
    inp, chans, ds = [[("conv", in_channels, ch)]], [ch], 1
    for level, mult in enumerate(mults):
        for _ in range(nrb):
            blk = [("res", ch, int(mult * mc), None)]
            ch = int(mult * mc)
            if ds in attn_ds:
                blk.append(("attn", ch, heads(ch, nh)))
            inp.append(blk)
            chans.append(ch)
        if level != len(mults) - 1:
            inp.append([("res", ch, ch, "down")])
            chans.append(ch)
            ds *= 2
    mid = [("res", ch, ch, None), ("attn", ch, heads(ch, nh)), ("res", ch, ch, None)]
    out = []
    for level, mult in list(enumerate(mults))[::-1]:
        for i in range(nrb + 1):
            ich = chans.pop()
            blk = [("res", ch + ich, int(mc * mult), None)]
            ch = int(mc * mult)
            if ds in attn_ds:
                blk.append(("attn", ch, heads(ch, nhu)))
            if level and i == nrb:
                blk.append(("res", ch, ch, "up"))
                ds //= 2
            out.append(blk)
    return inp, mid, out


def text_emb(P, o, tokens, mask):
    """text2im_model.py:79-91 + xf.py."""
    W, heads = o["xf_width"], o["xf_heads"]
    x = P["token_embedding.embedding_table"][tokens.long()] + P["positional_embedding"][None]
    if o["xf_padding"]:
        x = torch.where(mask.bool()[..., None], x, P["padding_embedding"][None])
    bs, n_ctx, _ = x.shape
    scale = 1 / math.sqrt(math.sqrt(W // heads))
    for i in range(o["xf_layers"]):
        b = "transformer.resblocks.%d." % i
        h = F.layer_norm(x, (W,), P[b + "ln_1.gamma"], P[b + "ln_1.beta"], 1e-7)
        qkv = linear(P, b + "attn.c_qkv.", h).reshape(bs, n_ctx, heads, -1)          # xf.py:86
        q, k, v = torch.chunk(qkv, 3, dim=-1)
        w = torch.matmul((q * scale).permute(0, 2, 1, 3), (k * scale).permute(0, 2, 3, 1))
        w = torch.softmax(w.float(), dim=-1)
        a = torch.matmul(w, v.permute(0, 2, 1, 3)).permute(0, 2, 1, 3).reshape(bs, n_ctx, -1)
        x = x + linear(P, b + "attn.c_proj.", a)
        h = F.layer_norm(x, (W,), P[b + "ln_2.gamma"], P[b + "ln_2.beta"], 1e-7)
        x = x + linear(P, b + "mlp.c_proj.", F.gelu(linear(P, b + "mlp.c_fc.", h), approximate="tanh"))
    if o["xf_final_ln"]:
        x = F.layer_norm(x, (W,), P["final_ln.gamma"], P["final_ln.beta"], 1e-7)
    return linear(P, "transformer_proj.", x[:, -1]), x.permute(0, 2, 1)


def resize_bilinear_legacy(x, size):
    """ops.ResizeBilinear(size, align_corners=False), half_pixel_centers=False."""
    n, c, H, W = x.shape
    OH, OW = size
    fy = torch.arange(OH, dtype=torch.float32) * (np.float32(H) / np.float32(OH))
    fx = torch.arange(OW, dtype=torch.float32) * (np.float32(W) / np.float32(OW))
    y0, x0 = fy.floor().long(), fx.floor().long()
    y1, x1 = (y0 + 1).clamp(max=H - 1), (x0 + 1).clamp(max=W - 1)
    wy, wx = (fy - y0)[:, None], (fx - x0)[None, :]
    g = lambda yy, xx: x[:, :, yy][:, :, :, xx]
    top = g(y0, x0) + (g(y0, x1) - g(y0, x0)) * wx
    bot = g(y1, x0) + (g(y1, x1) - g(y1, x0)) * wx
    return top + (bot - top) * wy


def unet_forward(P, o, x, timesteps, tokens, mask, low_res=None, taps=None):
    """Text2ImUNet.construct (text2im_model.py:93-121) / SuperResText2ImUNet.construct (:190-220)."""
    P = {k: (v if isinstance(v, torch.Tensor) else torch.from_numpy(np.asarray(v))).float() for k, v in P.items()}
    mc = o["num_channels"]
    if low_res is not None:
        x = torch.cat((x, resize_bilinear_legacy(low_res, (o["image_size"], o["image_size"]))), dim=1)
    inp, mid, out = structure(o, x.shape[1])
    emb = F.linear(timestep_embedding(timesteps, mc), P["time_embed.0.weight"], P["time_embed.0.bias"])
    emb = F.linear(F.silu(emb), P["time_embed.2.weight"], P["time_embed.2.bias"])
    xf_proj, xf_out = text_emb(P, o, tokens, mask)
    emb = emb + xf_proj

    def run(pre, layer, h):
        if layer[0] == "conv":
            return conv(P, pre, h, 1)
        if layer[0] == "res":
            return res_block(P, pre, h, emb, layer[1], layer[2], layer[3])
        return attention_block(P, pre, h, layer[2], xf_out)

    h, hs = x, []
    for bi, blk in enumerate(inp):
        for li, layer in enumerate(blk):
            h = run("input_blocks.%d.%d." % (bi, li), layer, h)
        hs.append(h)
        if taps is not None:
            taps["input_blocks.%d" % bi] = h
    for li, layer in enumerate(mid):
        h = run("middle_block.%d." % li, layer, h)
    if taps is not None:
        taps["middle_block"] = h
    for bi, blk in enumerate(out):
        h = torch.cat((h, hs.pop()), dim=1)
        for li, layer in enumerate(blk):
            h = run("output_blocks.%d.%d." % (bi, li), layer, h)
        if taps is not None:
            taps["output_blocks.%d" % bi] = h
    h = gn32(P, "out.0.", h, True)
    return conv(P, "out2.", h, 1)


# ---------------------------------------------------------------------------------------------------------
# samplers
# ---------------------------------------------------------------------------------------------------------
def p_mean_variance(T, model_output, x, i):
    """gaussian_diffusion.py:158-181 (clip_denoised; eps recomputed from the clipped x_start)."""
    eps, var = torch.chunk(model_output, 2, dim=1)
    frac = (var + 1) / 2
    logvar = frac * T.log_betas[i] + (1 - frac) * T.post_logvar[i]
    x0 = (T.sqrt_recip[i] * x - T.sqrt_recipm1[i] * eps).clamp(-1, 1)
    mean = T.coef1[i] * x0 + T.coef2[i] * x
    eps2 = (T.sqrt_recip[i] * x - x0) / T.sqrt_recipm1[i]
    return mean, logvar, x0, eps2


def guided_output(P, o, x, t_mapped, token, mask, random_token, random_mask, scale):
    """SamplingWithGuidance.construct (guider.py:85-90): CombinePrompt + model + Guider."""
    pics = x.shape[0] // 2
    half = x[:pics]
    tokens = torch.cat((token[:pics], random_token[None].expand(pics, -1)))
    masks = torch.cat((mask[:pics], random_mask[None].expand(pics, -1)))
    out = unet_forward(P, o, torch.cat((half, half)), t_mapped.expand(2 * pics), tokens, masks)
    eps, rest = out[:, :3], out[:, 3:]
    ce, ue = torch.chunk(eps, 2, dim=0)
    he = ue + scale * (ce - ue)
    return torch.cat((torch.cat((he, he)), rest), dim=1)


def p_sample_loop(P, o, token, mask, x_T, scale, random_tokens, step_noise, record=None):
    """main_funcs.py:20-41 + GenerativePSampleDiffusionModel (gaussian_diffusion.py:35-47).  random_tokens[i] and
    step_noise[i] (full 2*pics batch) are the draws of timestep i."""
    T = Tables(o["noise_schedule"], o["diffusion_steps"], o["timestep_respacing"])
    tmap = torch.tensor(T.timestep_map, dtype=torch.float32)
    img = x_T.float()
    ones = torch.ones(o["text_ctx"], dtype=torch.int32)
    for i in range(T.T - 1, -1, -1):
        out = guided_output(P, o, img, tmap[i:i + 1], token, mask, random_tokens[i], ones, scale)
        mean, logvar, x0, _ = p_mean_variance(T, out, img, i)
        img = mean + (0.0 if i == 0 else 1.0) * torch.exp(0.5 * logvar) * step_noise[i]
        if record is not None:
            record.append(img.clone())
    return img


def ddim_loop(P, o, samples, token, mask, x_T, eta=0.0):
    """main_funcs.py:44-65 + DDimSampleDiffusionModel / DDimSample (gaussian_diffusion.py:66-103, eta = 0)."""
    T = Tables(o["noise_schedule"], o["diffusion_steps"], o["timestep_respacing"])
    tmap = torch.tensor(T.timestep_map, dtype=torch.float32)
    low = torch.round((samples.float() + 1) * 127.5) / 127.5 - 1     # superResPreprocess, :224-230
    img = x_T.float()
    for i in range(T.T - 1, -1, -1):
        out = unet_forward(P, o, img, tmap[i:i + 1].expand(img.shape[0]), token, mask, low_res=low)
        _, _, x0, eps = p_mean_variance(T, out, img, i)
        ab, abp = T.ac_t[i], T.ac_prev_t[i]
        sigma = eta * torch.sqrt((1 - abp) / (1 - ab)) * torch.sqrt((1 - ab) / abp)
        img = torch.sqrt(1 - abp - sigma * sigma) * eps + x0 * torch.sqrt(abp)
    return img


# ---------------------------------------------------------------------------------------------------------
# SRGAN x4 stage (model/srgan.py)
# ---------------------------------------------------------------------------------------------------------
def depth_to_space2(x):
    """ops.DepthToSpace(2) on NCHW: out[n, c, 2h+i, 2w+j] = x[n, (2i+j)*C + c, h, w] (block offset is the major part
    of the depth index, as MindSpore / TensorFlow document it; NOT torch's pixel_shuffle order)."""
    n, c4, h, w = x.shape
    c = c4 // 4
    return x.reshape(n, 2, 2, c, h, w).permute(0, 3, 4, 1, 5, 2).reshape(n, c, 2 * h, 2 * w)


def srgan_forward(P, x, upscale_factor=4):
    """Generator.construct (srgan.py:96-104) with ResidualBlock (:47-54) and SubpixelConvolutionLayer (:64-68);
    nn.BatchNorm2d in inference mode (moving statistics, eps 1e-5), nn.PReLU(C) with per-channel slopes `w`."""
    P = {k: (v if isinstance(v, torch.Tensor) else torch.from_numpy(np.asarray(v))).float() for k, v in P.items()}
    conv2d = lambda p, t, pad: F.conv2d(t, P[p + "weight"], P[p + "bias"], padding=pad)
    bn = lambda p, t: F.batch_norm(t, P[p + "moving_mean"], P[p + "moving_variance"], P[p + "gamma"], P[p + "beta"],
                                   False, 0.0, 1e-5)
    prelu = lambda p, t: F.prelu(t, P[p + "w"])
    c1 = prelu("conv1.1.", conv2d("conv1.0.", x.float(), 4))
    t = c1
    for i in range(16):
        p = "trunk.%d." % i
        u = prelu(p + "prelu.", bn(p + "bn1.", conv2d(p + "conv1.", t, 1)))
        t = bn(p + "bn2.", conv2d(p + "conv2.", u, 1)) + t
    out = c1 + prelu("conv2.1.", conv2d("conv2.0.", t, 1))
    for i in range(int(math.log(upscale_factor, 2))):
        p = "subpixel_conv.%d." % i
        out = prelu(p + "prelu.", depth_to_space2(conv2d(p + "conv.", out, 1)))
    return torch.tanh(conv2d("conv3.", out, 4))
