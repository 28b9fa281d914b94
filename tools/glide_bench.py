"""Times the Taichu-GLIDE cascade (BASELINE.json configs[4]): 64x64 base (60 guided ancestral steps) + 256x256
upsampler (27 DDIM steps) on synthetic prompts and random-init weights of the shipped architectures.

    python tools/glide_bench.py --batch 32 [--base-steps 60] [--no-graph]
"""
import argparse
import json
import os
import sys
import time

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from minddiffusion_b200 import _lib, glide  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--batch", type=int, default=32)
    ap.add_argument("--base-steps", type=str, default="60")
    ap.add_argument("--scale", type=float, default=5.0)
    ap.add_argument("--no-graph", action="store_true")
    ap.add_argument("--skip-up", action="store_true")
    ap.add_argument("--srgan", action="store_true", help="also run the SRGAN x4 stage (256 -> 1024) on the result")
    a = ap.parse_args()
    P = a.batch
    ob = glide.model_and_diffusion_defaults(timestep_respacing=a.base_steps)
    ou = glide.model_and_diffusion_upsample()
    dev = "cuda"
    t0 = time.time()
    dm = glide.init_diffusion_model(ob, a.scale, (2 * P, 3, 64, 64))
    sm = None if a.skip_up else glide.init_super_res_model(ou, (P, 3, 256, 256))
    print("models built in %.1f s" % (time.time() - t0), file=sys.stderr)
    for m in (dm, sm):
        if m is not None:
            m.model.use_cuda_graph = not a.no_graph
    g = torch.Generator(device=dev).manual_seed(0)
    tok = torch.randint(1, ob["n_vocab"] - 1, (2 * P, 128), generator=g, device=dev, dtype=torch.int32)
    mask = torch.zeros(2 * P, 128, dtype=torch.int32, device=dev)
    mask[:, :24] = 1
    res = {}

    def timed(fn, name, steps):
        fn()                                  # warm-up (captures graphs)
        torch.cuda.synchronize()
        n0 = _lib.launch_count()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        out = fn()
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1)
        res[name] = {"ms": ms, "ms_per_step": ms / steps, "steps": steps, "eager_launches": _lib.launch_count() - n0}
        return out

    samples = timed(lambda: glide.gaussian_p_sample_loop(dm, tok, mask, (2 * P, 3, 64, 64), dm.num_timesteps,
                                                         vocab_len=ob["n_vocab"])[:P], "base", dm.num_timesteps)
    assert torch.isfinite(samples).all()
    if sm is not None:
        tok2 = torch.randint(1, ou["n_vocab"] - 1, (P, 128), generator=g, device=dev, dtype=torch.int32)
        up = timed(lambda: glide.ddim_sample_loop(sm, (P, 3, 256, 256), samples, tok2, mask[:P], sm.num_timesteps),
                   "upsampler", sm.num_timesteps)
        assert torch.isfinite(up).all()
        if a.srgan:
            from minddiffusion_b200.srgan import SRGAN
            sr = SRGAN(4)
            chunk = 8                                   # 1024^2 x 64-channel fp32 activations: 268 MB per image
            big = timed(lambda: torch.cat([sr.sr_handle(up[i:i + chunk].clamp(-1, 1)) for i in range(0, P, chunk)]),
                        "srgan", 1)
            assert tuple(big.shape) == (P, 3, 1024, 1024) and torch.isfinite(big).all()
    total = sum(v["ms"] for v in res.values())
    print(json.dumps({"workload": "Taichu-GLIDE cascade, batch %d" % P, "images_per_sec": P / (total / 1e3),
                      "total_ms": total, **res}))


if __name__ == "__main__":
    main()
