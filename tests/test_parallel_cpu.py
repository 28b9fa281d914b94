"""CPU tier: the N > 1 host logic on the gloo backend (world_size 2 and 4, rendezvous on 127.0.0.1).

Covers prompt-batch sharding, the max-over-ranks timing reduction, and the CFG pair split: rank 2k evaluates the
unconditional half, rank 2k+1 the conditional half, they exchange eps and both apply the same combine, so both end
up with the latents the single-process (concatenated-batch) formula gives (plms.py:192-197)."""
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    return port


def _worker(rank, world, port, q):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from minddiffusion_b200.parallel import CfgPairSplit, max_over_ranks, shard_range
        # every prompt is owned by exactly one rank
        lo, hi = shard_range(10, rank, world)
        owned = torch.zeros(10)
        owned[lo:hi] = 1
        dist.all_reduce(owned)
        assert torch.equal(owned, torch.ones(10))
        assert max_over_ranks(float(rank)) == float(world - 1)
        # CFG pair split with a stand-in "UNet": eps = x * w(context)
        split = CfgPairSplit()
        g = torch.Generator().manual_seed(100 + split.pair)  # same latents / contexts on both ranks of a pair
        x = torch.randn(3, 4, 2, 2, generator=g)
        c, uc = torch.randn(3, 5, generator=g), torch.randn(3, 5, generator=g)
        unet = lambda xx, ctx: xx * ctx.sum(dim=1)[:, None, None, None]
        ctx = split.my_context(c, uc)
        assert torch.equal(ctx, c if rank % 2 else uc)
        both = split.exchange(unet(x, ctx))
        assert both.shape[0] == 6
        e_u, e_c = both[:3], both[3:]
        guided = e_u + 7.5 * (e_c - e_u)
        ref_u, ref_c = unet(x, uc), unet(x, c)
        assert torch.equal(e_u, ref_u) and torch.equal(e_c, ref_c)
        # both ranks of the pair hold bit-identical guided eps -> latents stay in lock step without a broadcast
        gathered = [torch.empty_like(guided) for _ in range(2)]
        dist.all_gather(gathered, guided, group=split.group)
        assert torch.equal(gathered[0], gathered[1])
        q.put((rank, "ok"))
    except Exception as e:  # pragma: no cover - reported to the parent
        q.put((rank, "FAIL %r" % (e,)))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world", [2, 4])
def test_sharding_and_cfg_pair_split_on_gloo(world):
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, q)) for r in range(world)]
    for p in procs:
        p.start()
    results = [q.get(timeout=120) for _ in range(world)]
    for p in procs:
        p.join(timeout=60)
    assert sorted(results) == [(r, "ok") for r in range(world)], results


def test_shard_range_edges():
    from minddiffusion_b200.parallel import shard_range
    assert [shard_range(7, r, 3) for r in range(3)] == [(0, 3), (3, 5), (5, 7)]
    assert [shard_range(2, r, 4) for r in range(4)] == [(0, 1), (1, 2), (2, 2), (2, 2)]  # empty shards allowed
    with pytest.raises(ValueError):
        shard_range(4, 4, 4)
