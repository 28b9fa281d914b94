"""Parameter inventory and seeded synthetic weights for the UNet (no checkpoints ship with the reference and
there is no network here: benchmarks and parity tests use random-init weights of the exact architecture).

A random-init reference UNet outputs exactly zero because `zero_module` zero-initialises every ResBlock's second
conv, every SpatialTransformer.proj_out and the final conv (openaimodel.py:162-165, 521-526; attention.py:223-231;
SURVEY.md F7).  `random_unet_params` therefore draws EVERY tensor, including those, from a seeded generator, with
a per-tensor scale chosen so activations stay O(1) through the 38 blocks.
"""
import numpy as np

from .unet import PREFIX, unet_structure


def unet_param_shapes(cfg, prefix=PREFIX):
    """{MindSpore parameter name: shape}, the names `load_param_into_net` matches (txt2img.py:52-63)."""
    inp, mid, out = unet_structure(cfg)
    mc = cfg["model_channels"]
    ted = 4 * mc
    ctx = cfg["context_dim"]
    linear = cfg.get("use_linear_in_transformer", False)
    depth = cfg.get("transformer_depth", 1)
    S = {prefix + "time_embed.0.weight": (ted, mc), prefix + "time_embed.0.bias": (ted,),
         prefix + "time_embed.2.weight": (ted, ted), prefix + "time_embed.2.bias": (ted,)}

    def conv(name, cout, cin, k):
        S[name + ".weight"] = (cout, cin, k, k)
        S[name + ".bias"] = (cout,)

    def dense(name, cout, cin, bias=True):
        S[name + ".weight"] = (cout, cin)
        if bias:
            S[name + ".bias"] = (cout,)

    def norm(name, c):
        S[name + ".gamma"] = (c,)
        S[name + ".beta"] = (c,)

    def add(pre, layer):
        kind = layer[0]
        if kind == "conv":
            conv(pre + "conv", layer[2], layer[1], 3)
        elif kind == "res":
            cin, cout = layer[1], layer[2]
            norm(pre + "in_layers_norm", cin)
            conv(pre + "in_layers_conv.conv", cout, cin, 3)
            dense(pre + "emb_layers.1", cout, ted)
            norm(pre + "out_layers_norm", cout)
            conv(pre + "out_layers_conv.conv", cout, cout, 3)
            if cin != cout:
                conv(pre + "skip_connection.conv", cout, cin, 1)
        elif kind == "st":
            ch, heads, dh = layer[1], layer[2], layer[3]
            inner = heads * dh
            norm(pre + "norm", ch)
            if linear:
                dense(pre + "proj_in", inner, ch)
                dense(pre + "proj_out", ch, inner)
            else:
                conv(pre + "proj_in", inner, ch, 1)
                conv(pre + "proj_out", ch, inner, 1)
            for d in range(depth):
                tb = pre + "transformer_blocks.%d." % d
                for a, kd in (("attn1", inner), ("attn2", ctx)):
                    dense(tb + a + ".to_q", inner, inner, bias=False)
                    dense(tb + a + ".to_k", inner, kd, bias=False)
                    dense(tb + a + ".to_v", inner, kd, bias=False)
                    dense(tb + a + ".to_out.0", inner, inner)
                dense(tb + "ff.net.0.proj", 8 * inner, inner)
                dense(tb + "ff.net.2", inner, 4 * inner)
                for nm in ("norm1", "norm2", "norm3"):
                    norm(tb + nm, inner)
        elif kind == "down":
            conv(pre + "op.conv", layer[1], layer[1], 3)
        elif kind == "up":
            conv(pre + "conv.conv", layer[1], layer[1], 3)

    for bi, block in enumerate(inp):
        for li, layer in enumerate(block):
            add("%sinput_blocks.%d.%d." % (prefix, bi, li), layer)
    for li, layer in enumerate(mid):
        add("%smiddle_block.%d." % (prefix, li), layer)
    for bi, block in enumerate(out):
        for li, layer in enumerate(block):
            add("%soutput_blocks.%d.%d." % (prefix, bi, li), layer)
    norm(prefix + "out.0", mc)
    conv(prefix + "out.2.conv", cfg["out_channels"], mc, 3)
    return S


def random_unet_params(cfg, seed=0, prefix=PREFIX):
    """Seeded fp32 numpy parameters for every tensor of `unet_param_shapes(cfg)`."""
    rng = np.random.default_rng(seed)
    P = {}
    for name, shape in unet_param_shapes(cfg, prefix).items():
        if name.endswith(".gamma"):
            v = 1.0 + 0.1 * rng.standard_normal(shape)
        elif name.endswith(".beta") or name.endswith(".bias"):
            v = 0.05 * rng.standard_normal(shape)
        else:
            fan_in = int(np.prod(shape[1:]))
            gain = 1.0
            # residual branches end in these; keep them smaller so the trunk does not blow up over 38 blocks
            if "out_layers_conv" in name or "proj_out" in name or "to_out" in name or "ff.net.2" in name:
                gain = 0.5
            if "emb_layers" in name:
                gain = 0.5
            v = gain * rng.standard_normal(shape) / np.sqrt(fan_in)
        P[name] = v.astype(np.float32)
    return P


def merge_lora(params, lora_params, lora_rank=4, lora_alpha=4):
    """Fold LoRA deltas into the dense weights: W' = W + (alpha / rank) * B A.

    The reference swaps `nn.Dense` for MindPet's `tk.delta.LoRADense` in CrossAttention when `enable_lora` is set
    (wukong-huahua ldm/modules/attention.py:109-127, configs/v1-inference-chinese-lora.yaml:38-40; MindPet is a
    third-party dependency that is not vendored and not version-pinned by the reference).  MindPet's published
    forward is  y = x W^T + b + (alpha / rank) * (x A^T) B^T  with `mindpet_delta_lora_a` = A [rank, in] and
    `mindpet_delta_lora_b` = B [out, rank]; at inference (no dropout) that is a plain dense layer with W' above,
    so the hot path needs no LoRA kernel.  `lora_params` holds the `*.mindpet_delta_lora_a/_b` arrays (the LoRA
    checkpoint of txt2img.py --lora_ckpt_filepath); returns a new dict, `params` is not modified."""
    out = dict(params)
    scale = float(lora_alpha) / float(lora_rank)
    merged = 0
    for key, a in lora_params.items():
        if not key.endswith("mindpet_delta_lora_a"):
            continue
        stem = key[: -len("mindpet_delta_lora_a")]
        b = lora_params.get(stem + "mindpet_delta_lora_b")
        wkey = stem + "weight"
        if b is None or wkey not in out:
            raise KeyError("LoRA pair for %s is incomplete (need %smindpet_delta_lora_b and %s)" % (key, stem, wkey))
        a, b = np.asarray(a, dtype=np.float32), np.asarray(b, dtype=np.float32)
        if a.shape[0] != lora_rank or b.shape[1] != lora_rank:
            raise ValueError("%s: rank %d/%d does not match lora_rank %d" % (stem, a.shape[0], b.shape[1], lora_rank))
        w = np.asarray(out[wkey], dtype=np.float32)
        delta = scale * (b @ a)
        if delta.shape != w.reshape(w.shape[0], -1).shape:
            raise ValueError("%s: delta %s does not fit weight %s" % (stem, delta.shape, w.shape))
        out[wkey] = (w.reshape(delta.shape) + delta).reshape(w.shape)
        merged += 1
    if merged == 0:
        raise ValueError("no *.mindpet_delta_lora_a entries found in lora_params")
    return out
