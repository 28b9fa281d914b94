import sys, importlib.util
sys.path.insert(0,'/root/repo'); sys.path.insert(0,'/root/repo/tests')
import torch, numpy as np
from conftest import TINY_UNET, TINY_UNET_V1
from oracle import sd_unet
from minddiffusion_b200.unet import UNetModel
from minddiffusion_b200.weights import random_unet_params
def rel(a,b): return float((a.float().cpu()-b).abs().max()/b.abs().max())
for name,cfg,hw,B in (("tiny-sd2",TINY_UNET,16,2),("tiny-v1",TINY_UNET_V1,16,2),("wk-full",sd_unet.WUKONG_UNET,32,1),("sd2-full",sd_unet.SD2_UNET,32,1)):
    for seed in (0,1,2):
        P = random_unet_params(cfg, seed)
        Pt = {k: torch.from_numpy(v) for k,v in P.items()}
        rs = np.random.RandomState(seed+10)
        x = torch.from_numpy(rs.randn(B,4,hw,hw).astype(np.float32)); ctx = torch.from_numpy(rs.randn(B,77,cfg["context_dim"]).astype(np.float32))
        t = torch.tensor([981.0, 21.0][:B])
        ref = sd_unet.unet_forward(Pt, cfg, x, t, ctx)
        m = UNetModel(P, dtype=torch.bfloat16, **cfg); m.set_context(ctx.cuda())
        out = m.forward_nhwc(x.cuda(), t.cuda())[..., :4].permute(0,3,1,2)
        print(name, "seed", seed, "rel err %.3e" % rel(out, ref), flush=True)
        del m
        if "full" in name and seed >= 1: break
