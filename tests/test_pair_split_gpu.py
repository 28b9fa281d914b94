"""CFG pair split on two GPUs (-m gpu; skipped on a single-GPU box): rank 0 evaluates the unconditional half of the
guided batch, rank 1 the conditional half, each step the 4-channel eps halves cross NVLink in one NCCL all-gather
issued inside the captured step graph, and both ranks apply the same fp32 combine + update.  The two ranks' latents
must be bit-identical, and equal (to bf16 run-to-run rounding) to the single-GPU guided run."""
import os
import socket

import numpy as np
import pytest
import torch

from conftest import TINY_UNET

pytestmark = pytest.mark.gpu


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    return port


def _inputs():
    rs = np.random.RandomState(21)
    x_T = torch.from_numpy(rs.randn(2, 4, 16, 16).astype(np.float32))
    c = torch.from_numpy(rs.randn(2, 77, TINY_UNET["context_dim"]).astype(np.float32))
    uc = torch.from_numpy(rs.randn(2, 77, TINY_UNET["context_dim"]).astype(np.float32))
    return x_T, c, uc


def _sample(ldm, which, steps, dev, graph):
    from minddiffusion_b200 import samplers
    x_T, c, uc = _inputs()
    ldm.unet.use_cuda_graph = graph
    cls = {"ddim": samplers.DDIMSampler, "plms": samplers.PLMSSampler, "dpm": samplers.DPMSolverSampler}[which]
    return cls(ldm).sample(steps, 2, (4, 16, 16), c.to(dev), x_T=x_T.to(dev), verbose=False,
                           unconditional_guidance_scale=7.5, unconditional_conditioning=uc.to(dev))[0]


def _worker(rank, port, q):
    import torch.distributed as dist
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    torch.cuda.set_device(rank)
    dev = torch.device("cuda", rank)
    dist.init_process_group("nccl", rank=rank, world_size=2, device_id=dev)
    try:
        from minddiffusion_b200.ldm import LatentDiffusion
        from minddiffusion_b200.parallel import CfgPairSplit
        from minddiffusion_b200.unet import UNetModel
        from minddiffusion_b200.weights import random_unet_params
        unet = UNetModel(random_unet_params(TINY_UNET, 0), dtype=torch.bfloat16, device=dev, **TINY_UNET)
        ldm = LatentDiffusion(unet, timesteps=1000, linear_start=0.00085, linear_end=0.012)
        single = {w: _sample(ldm, w, 6, dev, graph=False) for w in ("ddim", "plms", "dpm")}   # both halves on one GPU
        ldm.cfg_split = CfgPairSplit()
        res = {}
        for which in ("ddim", "plms", "dpm"):
            for graph in (False, True):
                out = _sample(ldm, which, 6, dev, graph)
                both = [torch.empty_like(out) for _ in range(2)]
                dist.all_gather(both, out)
                rel = float((out - single[which]).abs().max() / single[which].abs().max())
                res["%s graph=%s" % (which, graph)] = (bool(torch.equal(both[0], both[1])), rel)
        q.put((rank, res))
    except Exception as e:  # pragma: no cover - reported to the parent
        import traceback
        q.put((rank, "FAIL %r\n%s" % (e, traceback.format_exc())))
    finally:
        dist.destroy_process_group()


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs two GPUs")
def test_cfg_pair_split_two_gpus_bit_identical_latents():
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, port, q), daemon=True) for r in range(2)]
    for p in procs:
        p.start()
    try:
        results = dict(q.get(timeout=240) for _ in range(2))
    finally:   # a rank stuck in a collective must not keep the test process alive
        for p in procs:
            p.join(timeout=20)
            if p.is_alive():
                p.kill()
    for rank, res in results.items():
        assert isinstance(res, dict), res
        for key, (identical, rel) in res.items():
            assert identical, (rank, key)          # the pair never drifts apart: no broadcast needed
            assert rel < 2e-3, (rank, key, rel)    # same result as the one-GPU guided run (bf16 rounding order only)
