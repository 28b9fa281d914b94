"""CFG pair split on two GPUs (-m gpu; skipped on a single-GPU box): rank 0 evaluates the unconditional half of the
guided batch, rank 1 the conditional half, each step the packed 4-channel eps halves cross NVLink in one NCCL
all-gather, and both ranks apply the same fp32 combine + update.  The two ranks' latents must be bit-identical (no
broadcast keeps them together) and equal, to bf16 run-to-run rounding, to the single-GPU guided run — for DDIM, PLMS
and DPM-Solver++, eager and with captured step graphs (tests/pair_split_worker.py is one rank)."""
import os
import subprocess
import sys

import pytest
import torch

from conftest import ROOT

pytestmark = pytest.mark.gpu


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs two GPUs")
@pytest.mark.parametrize("in_graph", ["0", "1"], ids=["exchange-between-graphs", "exchange-in-graph"])
def test_cfg_pair_split_two_gpus_bit_identical_latents(in_graph):
    env = dict(os.environ, MDX_SPLIT_IN_GRAPH=in_graph)
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr",
           "127.0.0.1", "--master-port", "29533", os.path.join(ROOT, "tests", "pair_split_worker.py")]
    try:
        r = subprocess.run(cmd, env=env, capture_output=True, text=True, timeout=200)
    except subprocess.TimeoutExpired as e:   # a rank stuck in a collective: report, do not hang the suite
        pytest.fail("pair-split ranks timed out:\n%s\n%s" % ((e.stdout or b"")[-2000:], (e.stderr or b"")[-2000:]))
    assert r.returncode == 0, (r.stdout[-2000:], r.stderr[-3000:])
    assert r.stdout.count("PAIR_SPLIT_OK") == 2, r.stdout[-2000:]
