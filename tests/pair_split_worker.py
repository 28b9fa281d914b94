"""One rank of the 2-GPU CFG pair-split check (launched by tests/test_pair_split_gpu.py through torch.distributed.run,
one process per GPU).  Prints one line per case and `PAIR_SPLIT_OK` at the end; any assertion ends the rank."""
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

from conftest import TINY_UNET  # noqa: E402


def inputs():
    rs = np.random.RandomState(21)
    x_T = torch.from_numpy(rs.randn(2, 4, 16, 16).astype(np.float32))
    c = torch.from_numpy(rs.randn(2, 77, TINY_UNET["context_dim"]).astype(np.float32))
    uc = torch.from_numpy(rs.randn(2, 77, TINY_UNET["context_dim"]).astype(np.float32))
    return x_T, c, uc


def sample(ldm, which, steps, dev, graph):
    from minddiffusion_b200 import samplers
    x_T, c, uc = inputs()
    ldm.unet.use_cuda_graph = graph
    cls = {"ddim": samplers.DDIMSampler, "plms": samplers.PLMSSampler, "dpm": samplers.DPMSolverSampler}[which]
    return cls(ldm).sample(steps, 2, (4, 16, 16), c.to(dev), x_T=x_T.to(dev), verbose=False,
                           unconditional_guidance_scale=7.5, unconditional_conditioning=uc.to(dev))[0]


def main():
    rank, local = int(os.environ["RANK"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    from minddiffusion_b200.ldm import LatentDiffusion
    from minddiffusion_b200.parallel import CfgPairSplit
    from minddiffusion_b200.unet import UNetModel
    from minddiffusion_b200.weights import random_unet_params
    unet = UNetModel(random_unet_params(TINY_UNET, 0), dtype=torch.bfloat16, device=dev, **TINY_UNET)
    ldm = LatentDiffusion(unet, timesteps=1000, linear_start=0.00085, linear_end=0.012)
    single = {w: sample(ldm, w, 5, dev, graph=False) for w in ("ddim", "plms", "dpm")}   # both halves on one GPU
    ldm.cfg_split = CfgPairSplit()
    for which in ("ddim", "plms", "dpm"):
        for graph in (False, True):
            out = sample(ldm, which, 5, dev, graph)
            both = [torch.empty_like(out) for _ in range(2)]
            dist.all_gather(both, out)
            rel = float((out - single[which]).abs().max() / single[which].abs().max())
            identical = bool(torch.equal(both[0], both[1]))
            print("rank %d %s graph=%s in_graph=%s: identical=%s rel_vs_single_gpu=%.2e" % (
                rank, which, graph, ldm.cfg_split.in_graph, identical, rel), flush=True)
            assert identical, "the two ranks of a pair drifted apart"
            assert rel < 2e-3, rel
    dist.barrier()
    torch.cuda.synchronize()
    print("rank %d PAIR_SPLIT_OK" % rank, flush=True)
    if ldm.cfg_split.in_graph:
        # graphs that captured an NCCL collective keep the communicator busy: destroy_process_group() did not return
        # on this stack (torch 2.11 / NCCL 2.28) while they were alive, so this rank leaves without the teardown
        sys.stdout.flush()
        os._exit(0)
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
