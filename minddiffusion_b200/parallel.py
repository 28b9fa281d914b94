"""Multi-GPU plumbing for the sampling path (one process per GPU, torch.distributed / NCCL over NVLink).

The reference samples on a single device (txt2img.py:198, `DEVICE_ID`); its only parallelism is data-parallel
training over HCCL.  The denoising path shards two ways, both exact:

  * prompt data parallelism — samples are independent (GroupNorm is per sample), so a prompt batch is split across
    ranks with NO collective inside the step loop (`shard_range`);
  * CFG pair split — the conditional and unconditional halves of the guided batch only meet in
    e_u + s (e_c - e_u) (plms.py:195-197, dpm_solver.py:321-325).  Rank 2k of a pair evaluates the unconditional
    half, rank 2k+1 the conditional half, and each step they exchange eps (`CfgPairSplit.exchange`, a 2-rank
    all-gather) and apply the identical fp32 combine + update, so the latents stay identical on both ranks without a
    broadcast.

Everything here also runs on CPU tensors with the gloo backend, which is how the host logic is tested.
"""
import torch
import torch.distributed as dist


def shard_range(n_items, rank, world):
    """Contiguous [start, stop) of `n_items` owned by `rank`; the first n_items % world ranks get one extra."""
    if world <= 0 or not (0 <= rank < world):
        raise ValueError("bad rank/world %r/%r" % (rank, world))
    base, extra = divmod(n_items, world)
    start = rank * base + min(rank, extra)
    return start, start + base + (1 if rank < extra else 0)


def max_over_ranks(value, device="cpu", group=None):
    """Max of a python float over all ranks (timing is reported as the slowest rank)."""
    if not dist.is_available() or not dist.is_initialized() or dist.get_world_size(group) == 1:
        return float(value)
    t = torch.tensor([float(value)], dtype=torch.float64, device=device)
    dist.all_reduce(t, op=dist.ReduceOp.MAX, group=group)
    return float(t.item())


class CfgPairSplit:
    """Splits classifier-free guidance across pairs of ranks (2k: unconditional, 2k+1: conditional)."""

    def __init__(self, rank=None, world=None):
        if not dist.is_initialized():
            raise RuntimeError("torch.distributed is not initialised")
        self.rank = dist.get_rank() if rank is None else rank
        self.world = dist.get_world_size() if world is None else world
        if self.world % 2:
            raise ValueError("CFG pair split needs an even number of ranks (got %d)" % self.world)
        self.pair = self.rank // 2
        self.is_cond = self.rank % 2 == 1
        if self.world == 2:
            self.group = dist.group.WORLD
        else:  # new_group is collective over the default group: every rank creates every pair's group, in order
            self.group = None
            for p in range(self.world // 2):
                g = dist.new_group(ranks=[2 * p, 2 * p + 1])
                if p == self.pair:
                    self.group = g
        self.n_pairs = self.world // 2
        # MDX_SPLIT_IN_GRAPH=1: the all-gather is captured inside the step graph (one graph per step); default: it is
        # issued between two captured segments of the step
        import os
        self.in_graph = os.environ.get("MDX_SPLIT_IN_GRAPH", "0") == "1"

    def my_context(self, cond, uncond):
        return cond if self.is_cond else uncond

    def pack(self, eps_local, packed, channels):
        """The first `channels` entries of every row of eps_local ([B, h, w, ld]: the tensor-core conv_out pads its 4
        output channels to 16 columns) into packed [B, h, w, channels] with one strided row copy, so that a quarter of
        the bytes cross NVLink."""
        ld = eps_local.shape[-1]
        if ld == channels:
            packed.copy_(eps_local)
        elif eps_local.is_cuda and (channels * eps_local.element_size()) % 16 == 0:
            from . import ops
            ops.copy_rows(eps_local.view(1, -1, ld)[:, :, :channels], packed.view(1, -1, channels))
        else:
            packed.copy_(eps_local[..., :channels])
        return packed

    def all_gather(self, packed, out):
        """out[2B, ...] = [unconditional rank's packed | conditional rank's packed]: one NCCL all-gather into a
        contiguous tensor (rank order = pair order = the concat order of plms.py:192-194)."""
        dist.all_gather_into_tensor(out, packed, group=self.group)
        return out

    def exchange(self, eps_local, out=None, channels=None):
        """eps_local: this rank's UNet output [B, ...].  Returns [2B, ...] with the unconditional half first — the
        row order the fused combine+update kernel expects.  channels: see pack()."""
        eps_local = eps_local.contiguous()
        if channels is not None and channels != eps_local.shape[-1]:
            eps_local = self.pack(eps_local, torch.empty(tuple(eps_local.shape[:-1]) + (channels,), dtype=eps_local.dtype,
                                                         device=eps_local.device), channels)
        if out is None:
            out = torch.empty((2 * eps_local.shape[0],) + tuple(eps_local.shape[1:]), dtype=eps_local.dtype,
                              device=eps_local.device)
        return self.all_gather(eps_local, out)
