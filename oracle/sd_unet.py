"""ORACLE (test infrastructure only — never imported by the product path).

CPU restatement, in plain torch fp32, of the reference's latent-diffusion UNet
(`/root/reference/vision/stablediffusionv2/ldm/modules/diffusionmodules/openaimodel.py`, `.../util.py`,
`ldm/modules/attention.py`; wukong-huahua's copies differ only in the constructor arguments).
Each function cites the reference lines it follows.  MindSpore semantics encoded explicitly: nn.Dense is
y = x W^T + b; nn.Conv2d(pad_mode='pad', padding=1) is symmetric zero padding; nn.GroupNorm / nn.LayerNorm use the
biased variance with eps inside the sqrt; ops.GeLU / nn.GELU() is the tanh approximation; ops.ResizeNearestNeighbor
maps dst -> floor(dst / 2).

Pinning status: the reference ships no tests or golden vectors and MindSpore cannot be installed
here (F9).  This restatement is pinned instead against the UNMODIFIED reference source executed in this container
on a torch-backed stand-in for the `mindspore` API (tests/golden/ms_shim.py + tests/golden/make_golden.py); the
resulting vectors are committed under tests/golden/ and checked by tests/test_oracle_golden.py.

Parameters are a flat dict {MindSpore parameter name: torch.float32 tensor}, names as `load_param_into_net` sees
them (prefix "model.diffusion_model." by default).
"""
import math

import torch
import torch.nn.functional as F

PREFIX = "model.diffusion_model."

SD2_UNET = dict(  # configs/v2-inference.yaml:21-37
    in_channels=4, out_channels=4, model_channels=320, attention_resolutions=[4, 2, 1], num_res_blocks=2,
    channel_mult=[1, 2, 4, 4], num_head_channels=64, num_heads=-1, use_spatial_transformer=True,
    use_linear_in_transformer=True, transformer_depth=1, context_dim=1024, legacy=False)

WUKONG_UNET = dict(  # wukong-huahua/configs/v1-inference-chinese.yaml:21-37
    in_channels=4, out_channels=4, model_channels=320, attention_resolutions=[4, 2, 1], num_res_blocks=2,
    channel_mult=[1, 2, 4, 4], num_head_channels=-1, num_heads=8, use_spatial_transformer=True,
    use_linear_in_transformer=False, transformer_depth=1, context_dim=768, legacy=False)


def timestep_embedding(timesteps, dim, max_period=10000):
    """util.py:111-131."""
    half = dim // 2
    freqs = torch.exp(-math.log(max_period) * torch.arange(0, half, dtype=torch.float32) / half)
    args = timesteps.to(torch.float32)[:, None] * freqs[None]
    emb = torch.cat((torch.cos(args), torch.sin(args)), dim=-1)
    if dim % 2:
        emb = torch.cat((emb, torch.zeros_like(emb[:, :1])), dim=-1)
    return emb


def unet_structure(cfg):
    """Follows the constructor loops openaimodel.py:350-519. Returns (input_blocks, middle, output_blocks), each
    a list of blocks, each block a list of layer tuples:
      ("conv", cin, cout) | ("res", cin, cout) | ("st", ch, heads, dim_head) | ("down", ch) | ("up", ch)."""
    mc = cfg["model_channels"]
    num_heads = cfg.get("num_heads", -1)
    nhc = cfg.get("num_head_channels", -1)
    legacy = cfg.get("legacy", True)
    attn_res = list(cfg["attention_resolutions"])
    mult_list = list(cfg["channel_mult"])
    nrb = cfg["num_res_blocks"]

    def heads_for(ch):  # openaimodel.py:375-382
        nonlocal num_heads
        if nhc == -1:
            dim_head = ch // num_heads
        else:
            num_heads = ch // nhc
            dim_head = nhc
        if legacy:
            dim_head = ch // num_heads  # use_spatial_transformer is True on every shipped config
        return num_heads, dim_head

    input_blocks = [[("conv", cfg["in_channels"], mc)]]
    chans = [mc]
    ch, ds = mc, 1
    for level, mult in enumerate(mult_list):
        for _ in range(nrb):
            layers = [("res", ch, mult * mc)]
            ch = mult * mc
            if ds in attn_res:
                h, dh = heads_for(ch)
                layers.append(("st", ch, h, dh))
            input_blocks.append(layers)
            chans.append(ch)
        if level != len(mult_list) - 1:
            input_blocks.append([("down", ch)])
            chans.append(ch)
            ds *= 2
    h, dh = heads_for(ch)
    middle = [("res", ch, ch), ("st", ch, h, dh), ("res", ch, ch)]
    output_blocks = []
    for level, mult in list(enumerate(mult_list))[::-1]:
        for i in range(nrb + 1):
            ich = chans.pop()
            layers = [("res", ch + ich, mc * mult)]
            ch = mc * mult
            if ds in attn_res:
                h, dh = heads_for(ch)
                layers.append(("st", ch, h, dh))
            if level and i == nrb:
                layers.append(("up", ch))
                ds //= 2
            output_blocks.append(layers)
    return input_blocks, middle, output_blocks


def _gn(x, P, name, eps):
    return F.group_norm(x, 32, P[name + ".gamma"], P[name + ".beta"], eps)


def res_block(P, pre, x, emb):
    """ResBlock.construct, non-updown / non-scale-shift branch, openaimodel.py:176-205."""
    h = F.silu(_gn(x, P, pre + "in_layers_norm", 1e-5))  # GroupNorm32 default eps, util.py:87-108
    h = F.conv2d(h, P[pre + "in_layers_conv.conv.weight"], P[pre + "in_layers_conv.conv.bias"], padding=1)
    emb_out = F.linear(F.silu(emb), P[pre + "emb_layers.1.weight"], P[pre + "emb_layers.1.bias"])
    h = h + emb_out[:, :, None, None]
    h = F.silu(_gn(h, P, pre + "out_layers_norm", 1e-5))
    h = F.conv2d(h, P[pre + "out_layers_conv.conv.weight"], P[pre + "out_layers_conv.conv.bias"], padding=1)
    if pre + "skip_connection.conv.weight" in P:
        x = F.conv2d(x, P[pre + "skip_connection.conv.weight"], P[pre + "skip_connection.conv.bias"])
    return x + h


def cross_attention(P, pre, x, context, heads):
    """CrossAttention.construct, attention.py:117-166 (mask path unused by the samplers)."""
    q = F.linear(x, P[pre + "to_q.weight"])
    ctx = x if context is None else context
    k = F.linear(ctx, P[pre + "to_k.weight"])
    v = F.linear(ctx, P[pre + "to_v.weight"])
    b, n, inner = q.shape
    d = inner // heads

    def split(t):
        return t.reshape(b, t.shape[1], heads, d).permute(0, 2, 1, 3).reshape(b * heads, t.shape[1], d)

    q, k, v = split(q), split(k), split(v)
    sim = torch.matmul(q, k.transpose(1, 2)) * (d ** -0.5)
    attn = torch.softmax(sim, dim=-1)
    out = torch.matmul(attn, v)
    out = out.reshape(b, heads, n, d).permute(0, 2, 1, 3).reshape(b, n, heads * d)
    return F.linear(out, P[pre + "to_out.0.weight"], P[pre + "to_out.0.bias"])


def feed_forward(P, pre, x):
    """FeedForward(glu=True) / GEGLU, attention.py:41-70."""
    proj = F.linear(x, P[pre + "net.0.proj.weight"], P[pre + "net.0.proj.bias"])
    a, gate = proj.chunk(2, dim=-1)
    h = a * F.gelu(gate, approximate="tanh")
    return F.linear(h, P[pre + "net.2.weight"], P[pre + "net.2.bias"])


def basic_transformer_block(P, pre, x, context, heads):
    """attention.py:181-185; LayerNorm eps 1e-5 (attention.py:176-178)."""
    c = x.shape[-1]

    def ln(t, name):
        return F.layer_norm(t, (c,), P[pre + name + ".gamma"], P[pre + name + ".beta"], 1e-5)

    x = cross_attention(P, pre + "attn1.", ln(x, "norm1"), None, heads) + x
    x = cross_attention(P, pre + "attn2.", ln(x, "norm2"), context, heads) + x
    x = feed_forward(P, pre + "ff.", ln(x, "norm3")) + x
    return x


def spatial_transformer(P, pre, x, context, heads, use_linear, depth=1):
    """SpatialTransformer.construct, attention.py:237-256; Normalize eps 1e-6 (attention.py:83-84)."""
    b, c, h, w = x.shape
    x_in = x
    x = _gn(x, P, pre + "norm", 1e-6)
    if not use_linear:
        x = F.conv2d(x, P[pre + "proj_in.weight"], P[pre + "proj_in.bias"])
    x = x.reshape(b, c, h * w).permute(0, 2, 1)
    if use_linear:
        x = F.linear(x, P[pre + "proj_in.weight"], P[pre + "proj_in.bias"])
    for d in range(depth):
        x = basic_transformer_block(P, pre + "transformer_blocks.%d." % d, x, context, heads)
    if use_linear:
        x = F.linear(x, P[pre + "proj_out.weight"], P[pre + "proj_out.bias"])
    x = x.reshape(b, h, w, c).permute(0, 3, 1, 2)
    if not use_linear:
        x = F.conv2d(x, P[pre + "proj_out.weight"], P[pre + "proj_out.bias"])
    return x + x_in


def _run_layer(P, cfg, pre, layer, h, emb, context):
    kind = layer[0]
    if kind == "conv":
        return F.conv2d(h, P[pre + "conv.weight"], P[pre + "conv.bias"], padding=1)
    if kind == "res":
        return res_block(P, pre, h, emb)
    if kind == "st":
        return spatial_transformer(P, pre, h, context, layer[2], cfg.get("use_linear_in_transformer", False),
                                   cfg.get("transformer_depth", 1))
    if kind == "down":  # Downsample.construct, openaimodel.py:80-88
        return F.conv2d(h, P[pre + "op.conv.weight"], P[pre + "op.conv.bias"], stride=2, padding=1)
    if kind == "up":  # Upsample.construct, openaimodel.py:53-60
        h = F.interpolate(h, scale_factor=2, mode="nearest")
        return F.conv2d(h, P[pre + "conv.conv.weight"], P[pre + "conv.conv.bias"], padding=1)
    raise ValueError(kind)


def unet_forward(P, cfg, x, timesteps, context, prefix=PREFIX, taps=None):
    """UNetModel.construct, openaimodel.py:536-576. x: [N,C,H,W] fp32, timesteps: [N], context: [N,L,D]."""
    ib, mid, ob = unet_structure(cfg)
    t_emb = timestep_embedding(timesteps, cfg["model_channels"])
    emb = F.linear(t_emb, P[prefix + "time_embed.0.weight"], P[prefix + "time_embed.0.bias"])
    emb = F.linear(F.silu(emb), P[prefix + "time_embed.2.weight"], P[prefix + "time_embed.2.bias"])
    hs = []
    h = x
    for bi, block in enumerate(ib):
        for li, layer in enumerate(block):
            h = _run_layer(P, cfg, "%sinput_blocks.%d.%d." % (prefix, bi, li), layer, h, emb, context)
        hs.append(h)
        if taps is not None:
            taps["input_blocks.%d" % bi] = h
    for li, layer in enumerate(mid):
        h = _run_layer(P, cfg, "%smiddle_block.%d." % (prefix, li), layer, h, emb, context)
    if taps is not None:
        taps["middle_block"] = h
    for bi, block in enumerate(ob):
        h = torch.cat((h, hs.pop()), dim=1)
        for li, layer in enumerate(block):
            h = _run_layer(P, cfg, "%soutput_blocks.%d.%d." % (prefix, bi, li), layer, h, emb, context)
        if taps is not None:
            taps["output_blocks.%d" % bi] = h
    h = F.silu(_gn(h, P, prefix + "out.0", 1e-5))
    return F.conv2d(h, P[prefix + "out.2.conv.weight"], P[prefix + "out.2.conv.bias"], padding=1)


def param_shapes(cfg, prefix=PREFIX):
    """{MindSpore parameter name: shape} for the UNet of `cfg` (what load_param_into_net would expect)."""
    ib, mid, ob = unet_structure(cfg)
    mc = cfg["model_channels"]
    ted = 4 * mc
    ctx = cfg["context_dim"]
    use_linear = cfg.get("use_linear_in_transformer", False)
    S = {}
    S[prefix + "time_embed.0.weight"] = (ted, mc)
    S[prefix + "time_embed.0.bias"] = (ted,)
    S[prefix + "time_embed.2.weight"] = (ted, ted)
    S[prefix + "time_embed.2.bias"] = (ted,)

    def add_layer(pre, layer):
        kind = layer[0]
        if kind == "conv":
            S[pre + "conv.weight"] = (layer[2], layer[1], 3, 3)
            S[pre + "conv.bias"] = (layer[2],)
        elif kind == "res":
            cin, cout = layer[1], layer[2]
            S[pre + "in_layers_norm.gamma"] = (cin,)
            S[pre + "in_layers_norm.beta"] = (cin,)
            S[pre + "in_layers_conv.conv.weight"] = (cout, cin, 3, 3)
            S[pre + "in_layers_conv.conv.bias"] = (cout,)
            S[pre + "emb_layers.1.weight"] = (cout, ted)
            S[pre + "emb_layers.1.bias"] = (cout,)
            S[pre + "out_layers_norm.gamma"] = (cout,)
            S[pre + "out_layers_norm.beta"] = (cout,)
            S[pre + "out_layers_conv.conv.weight"] = (cout, cout, 3, 3)
            S[pre + "out_layers_conv.conv.bias"] = (cout,)
            if cin != cout:
                S[pre + "skip_connection.conv.weight"] = (cout, cin, 1, 1)
                S[pre + "skip_connection.conv.bias"] = (cout,)
        elif kind == "st":
            ch, heads, dh = layer[1], layer[2], layer[3]
            inner = heads * dh
            S[pre + "norm.gamma"] = (ch,)
            S[pre + "norm.beta"] = (ch,)
            if use_linear:
                S[pre + "proj_in.weight"] = (inner, ch)
                S[pre + "proj_out.weight"] = (ch, inner)
            else:
                S[pre + "proj_in.weight"] = (inner, ch, 1, 1)
                S[pre + "proj_out.weight"] = (ch, inner, 1, 1)
            S[pre + "proj_in.bias"] = (inner,)
            S[pre + "proj_out.bias"] = (ch,)
            for d in range(cfg.get("transformer_depth", 1)):
                tb = pre + "transformer_blocks.%d." % d
                for a, kd in (("attn1", inner), ("attn2", ctx)):
                    S[tb + a + ".to_q.weight"] = (inner, inner)
                    S[tb + a + ".to_k.weight"] = (inner, kd)
                    S[tb + a + ".to_v.weight"] = (inner, kd)
                    S[tb + a + ".to_out.0.weight"] = (inner, inner)
                    S[tb + a + ".to_out.0.bias"] = (inner,)
                S[tb + "ff.net.0.proj.weight"] = (8 * inner, inner)
                S[tb + "ff.net.0.proj.bias"] = (8 * inner,)
                S[tb + "ff.net.2.weight"] = (inner, 4 * inner)
                S[tb + "ff.net.2.bias"] = (inner,)
                for nm in ("norm1", "norm2", "norm3"):
                    S[tb + nm + ".gamma"] = (inner,)
                    S[tb + nm + ".beta"] = (inner,)
        elif kind == "down":
            S[pre + "op.conv.weight"] = (layer[1], layer[1], 3, 3)
            S[pre + "op.conv.bias"] = (layer[1],)
        elif kind == "up":
            S[pre + "conv.conv.weight"] = (layer[1], layer[1], 3, 3)
            S[pre + "conv.conv.bias"] = (layer[1],)

    for bi, block in enumerate(ib):
        for li, layer in enumerate(block):
            add_layer("%sinput_blocks.%d.%d." % (prefix, bi, li), layer)
    for li, layer in enumerate(mid):
        add_layer("%smiddle_block.%d." % (prefix, li), layer)
    for bi, block in enumerate(ob):
        for li, layer in enumerate(block):
            add_layer("%soutput_blocks.%d.%d." % (prefix, bi, li), layer)
    S[prefix + "out.0.gamma"] = (mc,)
    S[prefix + "out.0.beta"] = (mc,)
    S[prefix + "out.2.conv.weight"] = (cfg["out_channels"], mc, 3, 3)
    S[prefix + "out.2.conv.bias"] = (cfg["out_channels"],)
    return S
