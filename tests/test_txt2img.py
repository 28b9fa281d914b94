"""The txt2img command line keeps the reference's argument surface; on a GPU the whole pipeline
(text tower -> guided sampler -> VAE decode -> PNG) runs end to end."""
import os

import numpy as np
import pytest


def test_argparse_surface_matches_reference_defaults():
    from minddiffusion_b200.txt2img import build_parser
    opt = build_parser().parse_args([])
    # vision/stablediffusionv2/txt2img.py:69-193
    want = dict(data_path="", prompt="A Van Gogh style oil painting of sunflower", output_path="output", skip_grid=False,
                skip_save=False, ddim_steps=50, fixed_code=False, ddim_eta=0.0, n_iter=2, H=512, W=512, n_samples=8,
                dpm_solver=False, n_rows=0, scale=9.0, from_file=None, config="configs/v2-inference.yaml",
                ckpt_path="models", ckpt_name="stablediffusionv2_512.ckpt", seed=42, precision="autocast")
    for k, v in want.items():
        assert getattr(opt, k) == v, k
    opt = build_parser().parse_args(["--dpm_solver", "--ddim_steps", "15", "--scale", "7.5", "--n_samples", "4",
                                     "--enable_lora", "False"])
    assert opt.dpm_solver and opt.ddim_steps == 15 and opt.scale == 7.5 and opt.n_samples == 4 and not opt.enable_lora


def test_prompt_token_stand_in_is_deterministic_and_well_formed():
    from minddiffusion_b200.txt2img import prompts_to_tokens
    with pytest.raises(RuntimeError, match="no token ids"):       # never invents ids silently
        prompts_to_tokens(["a cat"])
    a = prompts_to_tokens(["a cat", "", "a cat"], hashed=True)
    assert a.shape == (3, 77) and np.array_equal(a[0], a[2]) and not np.array_equal(a[0], a[1])
    assert a.min() >= 0 and a.max() < 49408 and a[0, 0] == 49406 and 49407 in a[0]
    table = {"a cat": np.arange(77)}
    assert np.array_equal(prompts_to_tokens(["a cat"], table)[0], np.arange(77))
    fn = lambda texts: np.full((len(texts), 77), 7)
    assert (prompts_to_tokens(["a dog"], table, tokenize_fn=fn) == 7).all()


def test_reference_yaml_is_understood(tmp_path):
    import yaml
    from minddiffusion_b200.txt2img import build_parser, load_config
    cfg = {"model": {"params": {"linear_start": 0.00085, "linear_end": 0.012, "timesteps": 1000, "scale_factor": 0.18215,
                                "unet_config": {"params": {"image_size": 32, "in_channels": 4, "out_channels": 4,
                                                           "model_channels": 320, "attention_resolutions": [4, 2, 1],
                                                           "num_res_blocks": 2, "channel_mult": [1, 2, 4, 4],
                                                           "num_heads": 8, "use_spatial_transformer": True,
                                                           "transformer_depth": 1, "context_dim": 768,
                                                           "use_checkpoint": True, "legacy": False, "use_fp16": True}},
                                "first_stage_config": {"params": {"ddconfig": {"double_z": True, "z_channels": 4,
                                                                                   "resolution": 256, "in_channels": 3,
                                                                                   "out_ch": 3, "ch": 128,
                                                                                   "ch_mult": [1, 2, 4, 4],
                                                                                   "num_res_blocks": 2,
                                                                                   "attn_resolutions": [],
                                                                                   "dropout": 0.0}}}}}}
    p = tmp_path / "v1.yaml"
    p.write_text(yaml.safe_dump(cfg))
    unet, dd, text, kw = load_config(build_parser().parse_args(["--config", str(p)]))
    assert unet["num_heads"] == 8 and unet["num_head_channels"] == -1 and not unet["use_linear_in_transformer"]
    assert "use_fp16" not in unet and dd["ch"] == 128 and text["width"] == 768 and kw["scale_factor"] == 0.18215


@pytest.mark.gpu
def test_txt2img_end_to_end_writes_images(tmp_path):
    from minddiffusion_b200.txt2img import main
    out = main(["--prompt", "a small test", "--H", "128", "--W", "128", "--n_samples", "2", "--n_iter", "1",
                "--ddim_steps", "2", "--scale", "7.5", "--output_path", str(tmp_path), "--config", "/nonexistent.yaml",
                "--hash_prompts"])
    assert len(out) == 1 and out[0].shape == (2, 3, 128, 128)
    assert np.isfinite(out[0]).all() and out[0].min() >= 0.0 and out[0].max() <= 1.0
    files = sorted(os.listdir(os.path.join(str(tmp_path), "samples")))
    assert files == ["00000.png", "00001.png"]
