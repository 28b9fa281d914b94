"""Reads the per-role (event, tile, clock) stamps CTA 0 of the persistent d <= 64 attention kernel writes in a
-DMDX_ATTN_TRACE build (MDX_LIB must point at one) and prints where each role's time goes per key tile."""
import ctypes
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch

from minddiffusion_b200 import _lib, ops

B, heads, N, d = 128, 5, 4096, 64
C = heads * d
torch.manual_seed(0)
q, k, v = (torch.randn(B, N, C, device="cuda").bfloat16() for _ in range(3))
for _ in range(3):
    ops.attention(q, k, v, heads, d ** -0.5)
torch.cuda.synchronize()
lib = _lib.load()
NR = 4096
buf = np.zeros(4 * NR * 2, dtype=np.uint32)
lib.mdx_attn_trace_dump.argtypes = [ctypes.c_void_p, ctypes.c_int]
rc = lib.mdx_attn_trace_dump(buf.ctypes.data, buf.size)
assert rc == 0, rc
buf = buf.reshape(4, NR, 2)
names = {0: "mma scores", 1: "softmax0", 2: "softmax1", 3: "mma P V"}
ev_names = {0: {1: "k_full ok", 2: "QK issued", 3: "p_full ok", 4: "v_full ok", 5: "PV issued"},
            1: {0: "loop top", 1: "s_full ok", 2: "S loaded", 3: "max done", 4: "o_full ok", 5: "exp done", 6: "P stored", 7: "rescale"},
            3: {3: "p,v_full ok", 5: "PV issued"}}
ev_names[2] = ev_names[1]
for role in range(4):
    rec = buf[role]
    n = int((rec[:, 1] != 0).sum())
    ev = (rec[:n, 0] >> 24).astype(int)
    tile = (rec[:n, 0] & 0xffffff).astype(int)
    clk = rec[:n, 1].astype(np.int64)
    dt = np.diff(clk) & 0xffffffff
    print("== %s: %d records" % (names[role], n))
    # steady state: tiles 8 .. 120 — time spent BEFORE each event (since the previous stamp of this role)
    sel = (tile[1:] >= 8) & (tile[1:] < 120)
    tot = 0.0
    for e in sorted(set(ev[1:][sel])):
        m = sel & (ev[1:] == e)
        print("   -> %-12s mean %7.1f  median %7.1f  p90 %7.1f  (n=%d)" % (ev_names[role].get(e, str(e)), dt[m].mean(), np.median(dt[m]), np.percentile(dt[m], 90), m.sum()))
        tot += dt[m].sum()
    ntiles = len(set(tile[1:][sel]))
    print("   cycles per own tile: %.1f over %d tiles" % (tot / max(ntiles, 1), ntiles))
    if role == 0:
        first = {}
        for e, t, c in zip(ev, tile, clk):
            first.setdefault((e, t), c)
        span = [first[(2, t + 1)] - first[(2, t)] for t in range(8, 119) if (2, t + 1) in first and (2, t) in first]
        print("   QK issue to QK issue: mean %.1f" % np.mean(span))
np.save(os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "gpurun_out", "attn_trace.npy"), buf)
