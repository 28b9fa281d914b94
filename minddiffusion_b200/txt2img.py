"""txt2img with the reference's command line (vision/stablediffusionv2/txt2img.py:66-193, wukong-huahua adds
--enable_lora / --lora_ckpt_filepath), running text tower -> sampler loop -> VAE decode on the sm_100a kernels.

    python -m minddiffusion_b200.txt2img --prompt "..." --H 512 --W 512 --n_samples 8 --ddim_steps 50 --scale 9.0
                                         [--dpm_solver] [--config configs/v2-inference.yaml] [--ckpt_path models]

Differences that follow from this build's scope:
  * checkpoints: MindSpore `.ckpt` files are protobuf and need MindSpore to read; this CLI loads an `.npz` holding the
    same parameter names (`model.diffusion_model...`, `first_stage_model...`, `cond_stage_model...`).  If the file does
    not exist it prints the reference's warning and continues with seeded random weights, as the reference continues
    with its random init (txt2img.py:52-63).
  * tokenisation: `minddiffusion_b200.tokenizer` (CLIP BPE for stablediffusionv2, WordPiece for wukong-huahua) with
    the reference's vocabulary files, which are looked up in $MDX_VOCAB_DIR / <repo>/vocab / a reference checkout;
    `--token_file` (int array [n_prompts, 77]) supplies ids directly.  Without either the CLI stops with an error;
    `--hash_prompts` is the explicit opt-in to a deterministic hash of the prompt (pipeline timing only: the images
    will not depict the prompt, and the CLI says so).
  * --config takes the reference YAML (read with `yaml`; OmegaConf is not needed) or defaults to the shipped values.
"""
import argparse
import hashlib
import os
import time

import numpy as np
import torch

from . import samplers
from .autoencoder import SD_DDCONFIG, AutoencoderKL, random_decoder_params
from .ldm import LatentDiffusion
from .text_encoder import SD2_TEXT, WUKONG_TEXT, FrozenCLIPEmbedder_ZH, random_text_params
from .unet import UNetModel
from .weights import random_unet_params

SD2_UNET = dict(in_channels=4, out_channels=4, model_channels=320, attention_resolutions=[4, 2, 1], num_res_blocks=2,
                channel_mult=[1, 2, 4, 4], num_head_channels=64, num_heads=-1, use_spatial_transformer=True,
                use_linear_in_transformer=True, transformer_depth=1, context_dim=1024, legacy=False)
WUKONG_UNET = dict(in_channels=4, out_channels=4, model_channels=320, attention_resolutions=[4, 2, 1], num_res_blocks=2,
                   channel_mult=[1, 2, 4, 4], num_head_channels=-1, num_heads=8, use_spatial_transformer=True,
                   use_linear_in_transformer=False, transformer_depth=1, context_dim=768, legacy=False)


def build_parser():
    p = argparse.ArgumentParser()
    p.add_argument("--data_path", type=str, nargs="?", default="", help="dir for prompts")
    p.add_argument("--prompt", type=str, nargs="?", default="A Van Gogh style oil painting of sunflower")
    p.add_argument("--output_path", type=str, nargs="?", default="output", help="dir to write results to")
    p.add_argument("--skip_grid", action="store_true")
    p.add_argument("--skip_save", action="store_true")
    p.add_argument("--ddim_steps", type=int, default=50, help="number of ddim sampling steps")
    p.add_argument("--fixed_code", action="store_true", help="use the same starting code across samples")
    p.add_argument("--ddim_eta", type=float, default=0.0, help="ddim eta (eta=0.0 is deterministic sampling)")
    p.add_argument("--n_iter", type=int, default=2, help="sample this often")
    p.add_argument("--H", type=int, default=512, help="image height, in pixel space")
    p.add_argument("--W", type=int, default=512, help="image width, in pixel space")
    p.add_argument("--n_samples", type=int, default=8, help="samples per prompt (batch size)")
    p.add_argument("--dpm_solver", action="store_true", help="use dpm_solver sampling")
    p.add_argument("--n_rows", type=int, default=0)
    p.add_argument("--scale", type=float, default=9.0, help="unconditional guidance scale")
    p.add_argument("--from-file", type=str, help="if specified, load prompts from this file")
    p.add_argument("--config", type=str, default="configs/v2-inference.yaml")
    p.add_argument("--ckpt_path", type=str, default="models")
    p.add_argument("--ckpt_name", type=str, default="stablediffusionv2_512.ckpt")
    p.add_argument("--seed", type=int, default=42)
    p.add_argument("--precision", type=str, choices=["full", "autocast"], default="autocast")
    p.add_argument("--enable_lora", default=False, type=lambda s: str(s).lower() in ("1", "true", "yes"))
    p.add_argument("--lora_ckpt_filepath", type=str, default="")
    p.add_argument("--lora_rank", type=int, default=4, help="unet_config.params.lora_rank of the reference YAML")
    p.add_argument("--lora_alpha", type=int, default=4, help="unet_config.params.lora_alpha of the reference YAML")
    # additions of this build
    p.add_argument("--ddim", action="store_true", help="use the DDIM sampler (eta-parameterised) instead of PLMS")
    p.add_argument("--token_file", type=str, default="", help=".npy int array [n_prompts, 77] of token ids")
    p.add_argument("--family", choices=["sd2", "wukong"], default="sd2", help="used when --config is not readable")
    p.add_argument("--hash_prompts", action="store_true",
                   help="no tokenizer vocabulary at hand: stand-in token ids hashed from the prompt (timing runs only)")
    return p


def load_config(opt):
    """(unet_cfg, ddconfig, text_cfg, ldm kwargs) from the reference YAML if readable, else the shipped defaults."""
    unet, dd = (SD2_UNET, SD_DDCONFIG) if opt.family == "sd2" else (WUKONG_UNET, SD_DDCONFIG)
    text = SD2_TEXT if opt.family == "sd2" else WUKONG_TEXT
    kw = dict(timesteps=1000, linear_start=0.00085, linear_end=0.012, scale_factor=0.18215)
    if os.path.exists(opt.config):
        import yaml
        prm = yaml.safe_load(open(opt.config))["model"]["params"]
        keep = set(SD2_UNET)
        unet = {k: v for k, v in prm["unet_config"]["params"].items() if k in keep}
        unet.setdefault("num_heads", -1)
        unet.setdefault("num_head_channels", -1)
        unet.setdefault("use_linear_in_transformer", False)
        dd = prm["first_stage_config"]["params"]["ddconfig"]
        kw = dict(timesteps=prm.get("timesteps", 1000), linear_start=prm["linear_start"], linear_end=prm["linear_end"],
                  scale_factor=prm.get("scale_factor", 1.0))
        text = SD2_TEXT if unet["context_dim"] == 1024 else WUKONG_TEXT
    return unet, dd, text, kw


def load_model(opt, device="cuda"):
    unet_cfg, dd, text_cfg, kw = load_config(opt)
    ckpt = os.path.join(opt.ckpt_path, opt.ckpt_name)
    print("Loading model from %s" % ckpt)
    npz = os.path.splitext(ckpt)[0] + ".npz"
    if os.path.exists(npz):
        params = dict(np.load(npz))
    else:
        print("!!!Warning!!!: %s doesn't exist" % ckpt)  # txt2img.py:60-61; the reference continues on random init
        params = {}
        params.update(random_unet_params(unet_cfg, opt.seed))
        params.update(random_decoder_params(dd, opt.seed))
        params.update(random_text_params(text_cfg, opt.seed))
    if opt.enable_lora:  # txt2img.py:60-63 of wukong-huahua loads the LoRA checkpoint on top of the base one
        from .weights import merge_lora
        lora_npz = os.path.splitext(opt.lora_ckpt_filepath)[0] + ".npz"
        if not os.path.exists(lora_npz):
            raise FileNotFoundError("--enable_lora needs %s (tools/ckpt_to_npz.py converts a MindSpore .ckpt)" % lora_npz)
        params = merge_lora(params, dict(np.load(lora_npz)), opt.lora_rank, opt.lora_alpha)
    dtype = torch.float32 if opt.precision == "full" else torch.bfloat16
    unet = UNetModel(params, dtype=dtype, device=device, **unet_cfg)
    unet.use_cuda_graph = True
    model = LatentDiffusion(unet, first_stage_model=AutoencoderKL(params, ddconfig=dd, dtype=dtype, device=device),
                            cond_stage_model=FrozenCLIPEmbedder_ZH(params, dtype=dtype, device=device, **text_cfg), **kw)
    return model


def prompts_to_tokens(prompts, token_table=None, context_length=77, vocab=49408, tokenize_fn=None, hashed=False):
    """Token ids [n, 77] of the prompts: from `token_table` (prompt -> ids) where it has the prompt, else from
    `tokenize_fn` (the family's tokenizer, encoders/modules.py:31-36).  hashed=True replaces a missing tokenizer by a
    deterministic stand-in (well-formed ids that do not encode the prompt)."""
    out = np.zeros((len(prompts), context_length), dtype=np.int64)
    for i, p in enumerate(prompts):
        if token_table is not None and p in token_table:
            out[i] = token_table[p]
            continue
        if tokenize_fn is not None:
            out[i] = tokenize_fn([p])[0]
            continue
        if not hashed:
            raise RuntimeError("no token ids for prompt %r: supply the tokenizer vocabulary ($MDX_VOCAB_DIR), a "
                               "--token_file, or pass --hash_prompts for a timing run" % p)
        h = hashlib.sha256(p.encode("utf-8")).digest()
        rs = np.random.RandomState(int.from_bytes(h[:4], "little"))
        n = min(context_length - 2, max(1, len(p.split())))
        out[i, 0] = vocab - 2                                  # <start_of_text>
        out[i, 1:1 + n] = rs.randint(1, vocab - 2, size=n)
        out[i, 1 + n] = vocab - 1                              # <end_of_text>
    return out


def family_tokenizer(cond_stage_model):
    """The tokenizer FrozenCLIPEmbedder_ZH pairs with its tower: CLIP BPE `tokenize` for stablediffusionv2
    (encoders/modules.py:28), WordPiece with [CLS]/[SEP] for wukong-huahua (its modules.py:28-50).  None when the
    vocabulary file cannot be found (the caller then needs --token_file for every prompt)."""
    from . import tokenizer
    zh = cond_stage_model.transformer.cfg.get("quick_gelu", False)      # the wukong tower's marker (WUKONG_TEXT)
    try:
        if zh:
            tk = tokenizer.WordpieceTokenizer()
            return lambda texts: tokenizer.tokenize_zh(texts, tokenizer=tk)
        tk = tokenizer.SimpleTokenizer()
        return lambda texts: tokenizer.tokenize(texts, truncate=True, tokenizer=tk)
    except FileNotFoundError as e:
        print("tokenizer vocabulary not found (%s)" % e)
        return None


def main(argv=None):
    opt = build_parser().parse_args(argv)
    torch.manual_seed(opt.seed)
    np.random.seed(opt.seed)
    model = load_model(opt)
    if opt.dpm_solver:
        sampler = samplers.DPMSolverSampler(model)
    elif opt.ddim:
        sampler = samplers.DDIMSampler(model)
    else:
        sampler = samplers.PLMSSampler(model)
    os.makedirs(opt.output_path, exist_ok=True)
    batch_size = opt.n_samples
    if not opt.data_path:
        assert opt.prompt is not None
        data = [batch_size * [opt.prompt]]
    else:
        path = os.path.join(opt.data_path, opt.prompt)
        print("reading prompts from %s" % path)
        with open(path, "r") as f:
            lines = f.read().splitlines()
        data = [batch_size * [p for p in lines]]
    token_table = None
    if opt.token_file:
        arr = np.load(opt.token_file)
        uniq = [p for p in dict.fromkeys(p for ps in data for p in ps)]
        token_table = {p: arr[i] for i, p in enumerate(uniq[:len(arr)])}
    sample_path = os.path.join(opt.output_path, "samples")
    os.makedirs(sample_path, exist_ok=True)
    base_count = len(os.listdir(sample_path))
    dev = model.unet.device
    start_code = None
    if opt.fixed_code:
        start_code = torch.randn((opt.n_samples, 4, opt.H // 8, opt.W // 8), device=dev)
    tokenize_fn = None if opt.hash_prompts else family_tokenizer(model.cond_stage_model)
    if opt.hash_prompts:
        print("!!!Warning!!!: --hash_prompts: token ids are hashed from the prompt text; the images will NOT depict it")
    encode = lambda ps: model.get_learned_conditioning(torch.from_numpy(
        prompts_to_tokens(ps, token_table, tokenize_fn=tokenize_fn, hashed=opt.hash_prompts)).to(dev))
    all_samples = []
    for prompts in data:
        for n in range(opt.n_iter):
            start_time = time.time()
            uc = encode(len(prompts) * [""]) if opt.scale != 1.0 else None
            c = encode(list(prompts))
            shape = [4, opt.H // 8, opt.W // 8]
            samples, _ = sampler.sample(S=opt.ddim_steps, conditioning=c, batch_size=len(prompts), shape=shape,
                                        verbose=False, unconditional_guidance_scale=opt.scale,
                                        unconditional_conditioning=uc, eta=opt.ddim_eta, x_T=start_code)
            x = model.decode_first_stage(samples)
            x = torch.clamp((x + 1.0) / 2.0, 0.0, 1.0).cpu().numpy()
            if not opt.skip_save:
                from PIL import Image
                for xs in x:
                    Image.fromarray((255.0 * xs.transpose(1, 2, 0)).astype(np.uint8)).save(
                        os.path.join(sample_path, "%05d.png" % base_count))
                    base_count += 1
            if not opt.skip_grid:
                all_samples.append(x)
            torch.cuda.synchronize()
            print("the infer time of a batch is %s" % (time.time() - start_time))
        print("Your samples are ready and waiting for you here: \n%s \n \nEnjoy." % opt.output_path)
    return all_samples


if __name__ == "__main__":
    main()
