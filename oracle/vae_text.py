"""ORACLE (test infrastructure only — never imported by the product path).

CPU restatement in torch fp32 of the two once-per-prompt networks of the stablediffusionv2 / wukong-huahua path:

  * `AutoencoderKL.decode` — ldm/models/autoencoder.py:65-68; `Decoder.construct` ldm/modules/diffusionmodules/
    model.py:408-440 with `ResnetBlock` (80-148, temb is None), `AttnBlock` (151-206), `Upsample` (31-52),
    `Normalize` = GroupNorm(32, eps 1e-6) (25-27), nonlinearity = x * sigmoid(x).
  * `TextEncoder.construct` — ldm/modules/encoders/text_encoder.py:114-153 (`MultiheadAttention` 25-64 with the
    triu(-inf, 1) mask of 131-134, `ResidualAttentionBlock` 80-98); SD2's "QuickGELU" is nn.GELU() (tanh form,
    67-69) and `ln_final` uses MindSpore's default eps 1e-7 (127); wukong's is x * sigmoid(1.702 x) with every
    LayerNorm at eps 1e-7 (wukong text_encoder.py:67-74,91,100,141).

Pinned like oracle/sd_unet.py: tests/golden/make_golden2.py executes the unmodified reference classes on the
torch-backed mindspore stand-in; tests/test_oracle_golden.py compares.
"""
import torch
import torch.nn.functional as F


def _gn(x, P, name):
    return F.group_norm(x, 32, P[name + ".gamma"], P[name + ".beta"], 1e-6)


def _swish(x):
    return x * torch.sigmoid(x)


def resnet_block(P, pre, x):
    """ResnetBlock.construct with temb=None (model.py:126-148)."""
    h = F.conv2d(_swish(_gn(x, P, pre + "norm1")), P[pre + "conv1.weight"], P[pre + "conv1.bias"], padding=1)
    h = F.conv2d(_swish(_gn(h, P, pre + "norm2")), P[pre + "conv2.weight"], P[pre + "conv2.bias"], padding=1)
    if pre + "nin_shortcut.weight" in P:
        x = F.conv2d(x, P[pre + "nin_shortcut.weight"], P[pre + "nin_shortcut.bias"])
    return x + h


def attn_block(P, pre, x):
    """AttnBlock.construct (model.py:186-206)."""
    h_ = _gn(x, P, pre + "norm")
    q = F.conv2d(h_, P[pre + "q.weight"], P[pre + "q.bias"])
    k = F.conv2d(h_, P[pre + "k.weight"], P[pre + "k.bias"])
    v = F.conv2d(h_, P[pre + "v.weight"], P[pre + "v.bias"])
    b, c, h, w = q.shape
    q = q.reshape(b, c, h * w).permute(0, 2, 1)
    k = k.reshape(b, c, h * w)
    w_ = torch.bmm(q, k) * (int(c) ** (-0.5))
    w_ = torch.softmax(w_, dim=2)
    v = v.reshape(b, c, h * w)
    h_ = torch.bmm(v, w_.permute(0, 2, 1)).reshape(b, c, h, w)
    return x + F.conv2d(h_, P[pre + "proj_out.weight"], P[pre + "proj_out.bias"])


def vae_decode(P, dd, z, prefix="first_stage_model."):
    """AutoencoderKL.decode: post_quant_conv then Decoder.construct. z: [B, z_channels, h, w] (already / scale_factor)."""
    d = prefix + "decoder."
    mult, nrb = list(dd["ch_mult"]), dd["num_res_blocks"]
    nres = len(mult)
    curr_res = dd["resolution"] // 2 ** (nres - 1)
    h = F.conv2d(z, P[prefix + "post_quant_conv.weight"], P[prefix + "post_quant_conv.bias"])
    h = F.conv2d(h, P[d + "conv_in.weight"], P[d + "conv_in.bias"], padding=1)
    h = resnet_block(P, d + "mid.block_1.", h)
    h = attn_block(P, d + "mid.attn_1.", h)
    h = resnet_block(P, d + "mid.block_2.", h)
    for i_level in reversed(range(nres)):
        for i_block in range(nrb + 1):
            h = resnet_block(P, d + "up.%d.block.%d." % (i_level, i_block), h)
            if curr_res in dd["attn_resolutions"]:
                h = attn_block(P, d + "up.%d.attn.%d." % (i_level, i_block), h)
        if i_level != 0:
            h = F.interpolate(h, scale_factor=2, mode="nearest")
            h = F.conv2d(h, P[d + "up.%d.upsample.conv.weight" % i_level], P[d + "up.%d.upsample.conv.bias" % i_level],
                         padding=1)
            curr_res *= 2
    h = _swish(_gn(h, P, d + "norm_out"))
    return F.conv2d(h, P[d + "conv_out.weight"], P[d + "conv_out.bias"], padding=1)


def text_encode(P, cfg, tokens, prefix="cond_stage_model.transformer."):
    """TextEncoder.construct (text_encoder.py:143-153). tokens: int [B, L]. cfg: width, layers, heads, quick_gelu,
    ln_eps (block LayerNorms), ln_final_eps."""
    width, heads = cfg["width"], cfg["heads"]
    B, L = tokens.shape
    x = P[prefix + "embedding_table"][tokens.reshape(-1).long()].reshape(B, L, width) + P[prefix + "positional_embedding"]
    mask = torch.triu(torch.full((L, L), float("-inf")), 1)
    d = width // heads
    for i in range(cfg["layers"]):
        b = prefix + "transformer_layer.resblocks.%d." % i
        h = F.layer_norm(x, (width,), P[b + "ln_1.gamma"], P[b + "ln_1.beta"], cfg["ln_eps"])
        qkv = F.linear(h, P[b + "attn.attn.in_proj.weight"], P[b + "attn.attn.in_proj.bias"])
        q, k, v = qkv.chunk(3, dim=-1)
        sp = lambda t: t.reshape(B, L, heads, d).permute(0, 2, 1, 3)
        s = torch.matmul(sp(q) * d ** -0.5, sp(k).transpose(-1, -2)) + mask
        o = torch.matmul(torch.softmax(s, -1), sp(v)).permute(0, 2, 1, 3).reshape(B, L, width)
        x = x + F.linear(o, P[b + "attn.attn.out_proj.weight"], P[b + "attn.attn.out_proj.bias"])
        h = F.layer_norm(x, (width,), P[b + "ln_2.gamma"], P[b + "ln_2.beta"], cfg["ln_eps"])
        h = F.linear(h, P[b + "c_fc.weight"], P[b + "c_fc.bias"])
        h = h * torch.sigmoid(1.702 * h) if cfg["quick_gelu"] else F.gelu(h, approximate="tanh")
        x = x + F.linear(h, P[b + "c_proj.weight"], P[b + "c_proj.bias"])
    return F.layer_norm(x, (width,), P[prefix + "ln_final.gamma"], P[prefix + "ln_final.beta"], cfg["ln_final_eps"])
