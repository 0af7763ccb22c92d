"""What bounds the claim kernel?  (GPU experiment, prints per-kernel microseconds.)

E1  new nodes vs none: route a fresh batch, then the same batch again (every node exists
    and is usable: pure lookups).
E2  table size at equal node count (2^23 .. 2^27 entries of 64 B): TLB reach / DRAM locality.
E3  claim decompositions 0 (warp per request) and 1 (warp per tile) for each of the above.

usage: python scripts/probe_claim.py
"""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "vllm-production-stack_b200"))
from b200_router import _lib, traces                      # noqa: E402
from b200_router.packing import pack_arrays              # noqa: E402
from b200_router.trie import GpuHashTrie                 # noqa: E402

dev = torch.device("cuda", 0)
B = 65536
NB = 6
tr = traces.SharedPrefixTrace(n_total=NB * B, batch=B, n_sys=64, sys_chars=2048, suffix_lo=256,
                              suffix_hi=1024, seed=7)
packed = [pack_arrays(*tr.batch_arrays(b)).to(dev) for b in range(NB)]
avail = 0xFF


def run(cap_log2, mode):
    _lib.check(_lib.lib().b200r_set_claim_by_tiles(mode))
    t = GpuHashTrie(cap_log2=cap_log2, device=dev)
    rows = []
    seq = 0
    _lib.prof_enable(True)
    for b in range(NB):
        for rep in range(2):                         # rep 1 = the same prompts again
            t.route(packed[b], avail, 1, seq)
            seq += B
            torch.cuda.synchronize()
            pr = _lib.prof_read()
            rows.append((b, rep, {k: v[0] * 1e3 for k, v in pr.items() if v[1]}))
    _lib.prof_enable(False)
    nodes = t.check_status(count_nodes=True)["num_nodes"]
    t.close()
    return rows, nodes


for cap in (23, 25, 27):
    for mode in (0, 1):
        rows, nodes = run(cap, mode)
        fresh = np.mean([r[2]["claim"] for r in rows if r[1] == 0 and r[0] >= 2])
        again = np.mean([r[2]["claim"] for r in rows if r[1] == 1 and r[0] >= 2])
        sel = np.mean([r[2]["select"] for r in rows if r[1] == 0 and r[0] >= 2])
        com = np.mean([r[2]["commit"] for r in rows if r[1] == 0 and r[0] >= 2])
        com2 = np.mean([r[2]["commit"] for r in rows if r[1] == 1 and r[0] >= 2])
        keys = np.mean([r[2].get("keys", 0.0) for r in rows if r[1] == 0 and r[0] >= 2])
        print(f"cap 2^{cap} ({(64 << cap) >> 20} MB) mode {mode}: claim fresh {fresh:6.1f} us, "
              f"same batch again {again:6.1f} us | select {sel:5.1f} commit fresh {com:5.1f} again {com2:5.1f} "
              f"keys {keys:5.1f} | nodes {nodes}", flush=True)
_lib.check(_lib.lib().b200r_set_claim_by_tiles(1))
