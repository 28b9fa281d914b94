"""Times AutoencoderKL.decode (SD VAE decoder, 64x64 latent -> 512x512) on random-init weights: python tools/vae_bench.py [B]"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from minddiffusion_b200.autoencoder import SD_DDCONFIG, AutoencoderKL, random_decoder_params

B = int(sys.argv[1]) if len(sys.argv) > 1 else 8
vae = AutoencoderKL(random_decoder_params(SD_DDCONFIG, 0), dtype=torch.bfloat16)
z = torch.randn(B, 4, 64, 64, device="cuda")
for _ in range(2):
    img = vae.decode(z)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(3):
    img = vae.decode(z)
e1.record()
torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / 3
# SD VAE decoder: 1.27 TFLOP (2*MAC) per 512x512 image
print("vae decode B=%d: %.2f ms  (%.2f ms/image, %.0f TFLOP/s)  finite=%s" % (
    B, ms, ms / B, B * 1.27e12 / ms / 1e9, bool(torch.isfinite(img).all())))
