"""FrozenCLIPEmbedder_ZH text tower on the sm_100a kernels (runs once per prompt, before the sampler loop).

Mirror of `TextEncoder` (vision/stablediffusionv2/ldm/modules/encoders/text_encoder.py:114-153, wukong-huahua's copy
for the SD-v1 shape) behind `FrozenCLIPEmbedder_ZH.construct` (encoders/modules.py:34-41): token + positional
embedding, `layers` x [LayerNorm -> fused in_proj -> causal multi-head attention -> out_proj (+residual);
LayerNorm -> c_fc + GELU -> c_proj (+residual)], ln_final.  Same parameter names
(`cond_stage_model.transformer.transformer_layer.resblocks.0.attn.attn.in_proj.weight`, ...).

Tokenisation (CLIP BPE / WordPiece, host string processing needing `ftfy` and vocabulary files) is out of scope
(SURVEY.md §2): `encode` takes token ids `[B, 77]`.
"""
import numpy as np
import torch

from . import ops

PREFIX = "cond_stage_model.transformer."

# encoders/modules.py:23-41 (SD2: OpenCLIP-H-like) and wukong-huahua encoders/modules.py:23-30 (CLIP-L-like)
SD2_TEXT = dict(context_length=77, vocab_size=49408, width=1024, layers=23, heads=16, quick_gelu=False, ln_eps=1e-5,
                ln_final_eps=1e-7)
WUKONG_TEXT = dict(context_length=77, vocab_size=49408, width=768, layers=12, heads=12, quick_gelu=True, ln_eps=1e-7,
                   ln_final_eps=1e-7)


def text_param_shapes(cfg, prefix=PREFIX):
    w = cfg["width"]
    S = {prefix + "embedding_table": (cfg["vocab_size"], w), prefix + "positional_embedding": (cfg["context_length"], w),
         prefix + "ln_final.gamma": (w,), prefix + "ln_final.beta": (w,)}
    for i in range(cfg["layers"]):
        b = prefix + "transformer_layer.resblocks.%d." % i
        S[b + "attn.attn.in_proj.weight"] = (3 * w, w)
        S[b + "attn.attn.in_proj.bias"] = (3 * w,)
        S[b + "attn.attn.out_proj.weight"] = (w, w)
        S[b + "attn.attn.out_proj.bias"] = (w,)
        for ln in ("ln_1", "ln_2"):
            S[b + ln + ".gamma"] = (w,)
            S[b + ln + ".beta"] = (w,)
        S[b + "c_fc.weight"] = (4 * w, w)
        S[b + "c_fc.bias"] = (4 * w,)
        S[b + "c_proj.weight"] = (w, 4 * w)
        S[b + "c_proj.bias"] = (w,)
    return S


def random_text_params(cfg, seed=0, prefix=PREFIX):
    rng = np.random.default_rng(seed)
    P = {}
    for name, shape in text_param_shapes(cfg, prefix).items():
        if name.endswith(".gamma"):
            v = 1.0 + 0.1 * rng.standard_normal(shape)
        elif name.endswith(".beta") or name.endswith(".bias"):
            v = 0.05 * rng.standard_normal(shape)
        elif name.endswith("embedding_table"):
            v = 0.5 * rng.standard_normal(shape)
        elif name.endswith("positional_embedding"):
            v = 0.1 * rng.standard_normal(shape)
        else:
            gain = 0.5 if ("out_proj" in name or "c_proj" in name) else 1.0
            v = gain * rng.standard_normal(shape) / np.sqrt(shape[1])
        P[name] = v.astype(np.float32)
    return P


class TextEncoder:
    def __init__(self, params, dtype=torch.bfloat16, device="cuda", prefix=PREFIX, **cfg):
        if not torch.cuda.is_available():
            raise RuntimeError("minddiffusion_b200 needs a CUDA device (sm_100a); there is no CPU path")
        self.cfg = dict(cfg)
        self.dtype, self.bf16 = dtype, dtype == torch.bfloat16
        self.device = torch.device(device)

        def g(n):
            key = prefix + n
            if key not in params:
                raise KeyError("missing parameter %s" % key)
            v = params[key]
            if isinstance(v, np.ndarray):
                v = torch.from_numpy(np.ascontiguousarray(v))
            return v.to(device=self.device, dtype=torch.float32).contiguous()

        w = lambda t: t.to(self.dtype).contiguous()
        self.table, self.pos = g("embedding_table"), g("positional_embedding")
        self.ln_final = (g("ln_final.gamma"), g("ln_final.beta"))
        self.blocks = []
        for i in range(cfg["layers"]):
            b = "transformer_layer.resblocks.%d." % i
            self.blocks.append(dict(
                ln1=(g(b + "ln_1.gamma"), g(b + "ln_1.beta")), ln2=(g(b + "ln_2.gamma"), g(b + "ln_2.beta")),
                qkv=(w(g(b + "attn.attn.in_proj.weight")), g(b + "attn.attn.in_proj.bias")),
                out=(w(g(b + "attn.attn.out_proj.weight")), g(b + "attn.attn.out_proj.bias")),
                fc=(w(g(b + "c_fc.weight")), g(b + "c_fc.bias")), proj=(w(g(b + "c_proj.weight")), g(b + "c_proj.bias"))))

    def _dense(self, x, w, b, residual=None, act=None, trunk=False):
        if self.bf16:
            return ops.gemm_tc([(x, 1)], w, bias=b, residual=residual, act=act,
                               out_dtype=torch.float32 if trunk else torch.bfloat16)
        return ops.conv_direct(x, w, bias=b, residual=residual, act=act)

    def __call__(self, tokens):
        """tokens: int [B, L] -> fp32 [B, L, width] (text_encoder.py:143-153)."""
        cfg, dt = self.cfg, self.dtype
        ids = tokens.to(device=self.device, dtype=torch.int32).contiguous()
        B, L = ids.shape
        if L != cfg["context_length"]:
            raise ValueError("expected %d tokens per prompt, got %d" % (cfg["context_length"], L))
        W, heads = cfg["width"], cfg["heads"]
        act = "quick_gelu" if cfg["quick_gelu"] else "gelu"
        x = ops.embedding(self.table, self.pos, ids)                       # fp32 token stream [B*L, W]
        for blk in self.blocks:
            h = ops.layernorm(x, *blk["ln1"], eps=cfg["ln_eps"], out_dtype=dt)
            qkv = self._dense(h, *blk["qkv"]).view(B, L, 3 * W)
            o = ops.attention(qkv[:, :, :W], qkv[:, :, W:2 * W], qkv[:, :, 2 * W:], heads, (W // heads) ** -0.5,
                              causal=True)
            x = self._dense(o.view(B * L, W), *blk["out"], residual=x, trunk=True)
            h = ops.layernorm(x, *blk["ln2"], eps=cfg["ln_eps"], out_dtype=dt)
            h = self._dense(h, *blk["fc"], act=act)
            x = self._dense(h, *blk["proj"], residual=x, trunk=True)
        return ops.layernorm(x, *self.ln_final, eps=cfg["ln_final_eps"], out_dtype=torch.float32).view(B, L, W)


class FrozenCLIPEmbedder_ZH:
    """encoders/modules.py:23-41: `encode(token ids [B, 77])` / `__call__(token ids)`; `tokenize(texts)` is the family's
    tokenizer (minddiffusion_b200.tokenizer) and `encode(list of str)` goes through it."""

    def __init__(self, params, dtype=torch.bfloat16, device="cuda", **cfg):
        self.transformer = TextEncoder(params, dtype=dtype, device=device, **(cfg or SD2_TEXT))
        self.max_length = self.transformer.cfg["context_length"]

    def tokenize(self, texts):
        from . import tokenizer
        if self.transformer.cfg.get("quick_gelu", False):      # wukong-huahua tower: WordPiece, [CLS] ... [SEP]
            return tokenizer.tokenize_zh(texts)
        return tokenizer.tokenize(texts, truncate=True)

    def encode(self, tokens):
        if isinstance(tokens, (str, list)) and (isinstance(tokens, str) or (tokens and isinstance(tokens[0], str))):
            import torch as _torch
            tokens = _torch.from_numpy(self.tokenize(tokens))   # FileNotFoundError if the vocabulary file is missing
        return self.transformer(tokens)

    __call__ = encode
