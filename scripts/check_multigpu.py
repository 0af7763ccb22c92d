#!/usr/bin/env python
"""Multi-GPU parity check (test infrastructure; launch with torchrun, one rank per GPU):

    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 \
        --master-port 29520 scripts/check_multigpu.py [part|nccl]

Routes a C3-shaped trace with availability churn through the multi-GPU router -- part:
partitioned table, real NVLink peer-memory exchange; nccl: replicated table, NCCL all-gather of
the chunk hashes -- and compares every rank's decisions, and the final table (union of the
partitions / every replica), with the sequential C oracle on the rank-major concatenation.
"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "vllm-production-stack_b200"))

import numpy as np
import torch
import torch.distributed as dist

from b200_router import traces
from b200_router.distributed import PartitionedPrefixRouter, ShardedPrefixRouter
from b200_router.packing import pack_arrays
from oracle import coracle


def main():
    mode = sys.argv[1] if len(sys.argv) > 1 else "part"
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    local = int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    nb, per = 6, 3000
    tr = traces.SharedPrefixTrace(n_total=nb * world * per, batch=per, n_sys=32, sys_chars=512,
                                  suffix_lo=100, suffix_hi=400, seed=31)
    shards = [[tr.batch_arrays(b * world + r) for r in range(world)] for b in range(nb)]
    max_chunks = max(int(((s[2] + 127) // 128).sum()) for row in shards for s in row)
    if mode == "part":
        router = PartitionedPrefixRouter(cap_log2=20, device=dev, max_requests=per,
                                         max_chunks=max_chunks)
    else:
        router = ShardedPrefixRouter(cap_log2=20, device=dev, max_requests=per, max_chunks=max_chunks)
    orc = coracle.TrieOracle()
    ok, seq = True, 0
    for b in range(nb):
        avail = 0xFF & ~(1 << 2) if b in (2, 3) else 0xFF
        ep, depth, matched = router.route(pack_arrays(*shards[b][rank], pin=False), avail, 17, seq)
        mine = (ep.cpu().numpy(), depth.cpu().numpy(), matched.cpu().numpy().view(np.uint64))
        off = 0
        for r in range(world):
            oep, odepth, omatched = orc.route_batch(*shards[b][r], avail, 17, seq + off)
            off += len(shards[b][r][1])
            if r == rank:
                ok &= bool((mine[0] == oep).all() and (mine[1] == odepth).all() and
                           (mine[2] == omatched).all())
        seq += off
    router.check_status()
    if mode == "part":                      # union of the partitions == the reference trie
        parts = [None] * world
        dist.all_gather_object(parts, router.dump())
        merged = {}
        for d in parts:
            for path, m in d.items():
                if path == ():
                    merged[path] = merged.get(path, 0) | m
                else:
                    ok &= path not in merged
                    merged[path] = m
        ok &= merged == orc.dump()
    else:
        ok &= router.dump() == orc.dump()
    flag = torch.tensor([1 if ok else 0], device=dev)
    dist.all_reduce(flag, op=dist.ReduceOp.MIN)
    if rank == 0:
        print("multi-GPU parity (%s, %d ranks): %s" % (mode, world, "OK" if flag.item() else "MISMATCH"))
    dist.destroy_process_group()
    sys.exit(0 if flag.item() else 1)


if __name__ == "__main__":
    main()
