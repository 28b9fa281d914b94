"""The once-per-prompt networks: VAE decoder and text tower.

CPU tier: the oracle (oracle/vae_text.py) against golden vectors from the executed reference
(tests/golden/make_golden2.py).  GPU tier: the CUDA path against the same goldens and against the oracle at a larger
size; fp32 check mode 1e-4 (decoder: 2e-4, ten chained blocks); bf16: text towers 1e-2, VAE decoder 2e-2 of the output
range (measured 0.9e-2 on the small decoder, 1.2e-2 on the shipped shape: 23 chained conv blocks in bf16 storage)."""
import os

import numpy as np
import pytest
import torch

from conftest import ROOT, parity_log

TINY_DD = dict(double_z=True, z_channels=4, resolution=32, in_channels=3, out_ch=3, ch=64, ch_mult=[1, 2, 2],
               num_res_blocks=1, attn_resolutions=[], dropout=0.0)
TINY_TEXT = {
    "sd2": dict(context_length=77, vocab_size=1000, width=128, layers=2, heads=2, quick_gelu=False, ln_eps=1e-5,
                ln_final_eps=1e-7),
    "wukong": dict(context_length=77, vocab_size=1000, width=128, layers=2, heads=2, quick_gelu=True, ln_eps=1e-7,
                   ln_final_eps=1e-7),
}


def golden(family):
    return dict(np.load(os.path.join(ROOT, "tests", "golden", "%s_vae_text_golden.npz" % family)))


def rel(a, b):
    a = a.detach().float().cpu().numpy() if isinstance(a, torch.Tensor) else np.asarray(a)
    b = b.detach().float().cpu().numpy() if isinstance(b, torch.Tensor) else np.asarray(b)
    return float(np.abs(a.astype(np.float64) - b).max() / (np.abs(b).max() + 1e-30))


def _z():
    return torch.from_numpy(np.random.RandomState(7).randn(2, 4, 8, 8).astype(np.float32))


@pytest.mark.parametrize("family", ["sd2", "wukong"])
def test_oracle_matches_executed_reference(family):
    from minddiffusion_b200.autoencoder import random_decoder_params
    from minddiffusion_b200.text_encoder import random_text_params
    from oracle import vae_text
    G = golden(family)
    P = {k: torch.from_numpy(v) for k, v in random_decoder_params(TINY_DD, 0).items()}
    assert rel(vae_text.vae_decode(P, TINY_DD, _z()), G["vae_image"]) < 1e-5
    tc = TINY_TEXT[family]
    PT = {k: torch.from_numpy(v) for k, v in random_text_params(tc, 0).items()}
    out = vae_text.text_encode(PT, tc, torch.from_numpy(G["text_tokens"]))
    assert rel(out, G["text_out"]) < 1e-5


def test_decoder_inventory():
    from minddiffusion_b200.autoencoder import SD_DDCONFIG, decoder_param_shapes, decoder_structure
    L = decoder_structure(SD_DDCONFIG)
    assert sum(l[0] == "res" for l in L) == 14 and sum(l[0] == "up" for l in L) == 3 and sum(l[0] == "attn" for l in L) == 1
    n = sum(int(np.prod(s)) for k, s in decoder_param_shapes(SD_DDCONFIG).items() if ".decoder." in k)
    assert 49.0e6 < n < 50.0e6  # the SD VAE decoder has ~49.5 M parameters


@pytest.mark.gpu
@pytest.mark.parametrize("family", ["sd2", "wukong"])
@pytest.mark.parametrize("dtype,tol_vae,tol_text", [(torch.float32, 2e-4, 1e-4), (torch.bfloat16, 2e-2, 1e-2)],
                         ids=["fp32", "bf16"])
def test_cuda_vae_and_text_against_goldens(family, dtype, tol_vae, tol_text):
    from minddiffusion_b200.autoencoder import AutoencoderKL, random_decoder_params
    from minddiffusion_b200.text_encoder import FrozenCLIPEmbedder_ZH, random_text_params
    G = golden(family)
    vae = AutoencoderKL(random_decoder_params(TINY_DD, 0), ddconfig=TINY_DD, dtype=dtype)
    img = vae.decode(_z().cuda())
    assert tuple(img.shape) == (2, 3, 32, 32)
    parity_log("vae tiny %s %s" % (family, dtype), rel(img, G["vae_image"]), tol_vae)
    assert rel(img, G["vae_image"]) < tol_vae
    tc = TINY_TEXT[family]
    te = FrozenCLIPEmbedder_ZH(random_text_params(tc, 0), dtype=dtype, **tc)
    out = te.encode(torch.from_numpy(G["text_tokens"]).cuda())
    assert tuple(out.shape) == (2, 77, 128) and out.dtype == torch.float32
    parity_log("text tower tiny %s %s" % (family, dtype), rel(out, G["text_out"]), tol_text)
    assert rel(out, G["text_out"]) < tol_text
    # prompts as strings go through the family's tokenizer (here on the small vocabularies committed with the goldens)
    from minddiffusion_b200 import tokenizer as T
    gold = os.path.join(ROOT, "tests", "golden")
    T._default.clear()
    if family == "sd2":
        T._default["bpe"] = T.SimpleTokenizer(os.path.join(gold, "mini_bpe.txt.gz"))
    else:
        T._default["zh"] = T.WordpieceTokenizer(os.path.join(gold, "mini_vocab_zh.txt"))
    try:
        ids = te.tokenize(["a painting of a cat", ""])
        assert ids.shape == (2, 77) and ids.max() < tc["vocab_size"] and (ids[1, 2:] == 0).all()
        from_text = te.encode(["a painting of a cat", ""])
        assert torch.equal(from_text, te.encode(torch.from_numpy(ids).cuda()))
    finally:
        T._default.clear()


@pytest.mark.gpu
def test_cuda_vae_decoder_reference_shape_bf16():
    """ch 128, mult (1,2,4,4), 2 blocks (the shipped ddconfig) on a 16x16 latent: 512-wide single-head attention over
    256 positions, 512 -> 256 -> 128 channel ResnetBlocks with nin_shortcut, three upsamples."""
    from minddiffusion_b200.autoencoder import SD_DDCONFIG, AutoencoderKL, random_decoder_params
    from oracle import vae_text
    P = random_decoder_params(SD_DDCONFIG, 1)
    z = torch.from_numpy(np.random.RandomState(3).randn(1, 4, 16, 16).astype(np.float32))
    ref = vae_text.vae_decode({k: torch.from_numpy(v) for k, v in P.items()}, SD_DDCONFIG, z)
    img = AutoencoderKL(P, dtype=torch.bfloat16).decode(z.cuda())
    assert tuple(img.shape) == (1, 3, 128, 128)
    parity_log("vae shipped-shape bf16", rel(img, ref), 2e-2)
    assert rel(img, ref) < 2e-2
