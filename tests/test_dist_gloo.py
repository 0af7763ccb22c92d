"""World-size-2 run of the sharded routing orchestration on CPUs (gloo).

The device steps are emulated (tests/fake_shard_ops.py); what is under test is
b200_router.distributed.ShardedPrefixRouter: shard bookkeeping, the three
all-gathers and the global-order semantics.  Result must equal routing the
rank-major concatenation of the shards sequentially (C oracle), and both
replicas must end up identical.
"""
import os
import pickle
import tempfile

import numpy as np
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from b200_router import traces
from b200_router.packing import pack_prompts
from oracle import coracle

WORLD = 2


def _shards():
    """[(avail mask, [prompts of rank 0], [prompts of rank 1])] -- uneven shard sizes,
    heavy sharing across ranks, availability churn, one empty shard."""
    out = []
    raw = traces.sharing_stress(n_batches=5, batch=60, n_endpoints=6, max_depth=5, alphabet=2, seed=3)
    for b, (avail_idx, prompts) in enumerate(raw):
        cut = [37, 0, 60, 20, 30][b]
        out.append((sum(1 << i for i in avail_idx), prompts[:cut], prompts[cut:]))
    return out


def _worker(rank, initfile, resfile):
    import sys
    here = os.path.dirname(os.path.abspath(__file__))
    sys.path.insert(0, here)
    from b200_router.distributed import ShardedPrefixRouter
    from fake_shard_ops import FakeShardOps
    dist.init_process_group("gloo", init_method="file://" + initfile, rank=rank, world_size=WORLD)
    router = ShardedPrefixRouter(ops=FakeShardOps())
    seq, results = 0, []
    for avail, p0, p1 in _shards():
        mine = (p0, p1)[rank]
        ep, depth, matched = router.route(pack_prompts(mine, pin=False), avail, 77, seq)
        seq += len(p0) + len(p1)
        results.append((ep.tolist(), depth.tolist(), matched.numpy().view(np.uint64).tolist()))
    with open(resfile + str(rank), "wb") as f:
        pickle.dump((results, router.dump()), f)
    dist.barrier()
    dist.destroy_process_group()


def test_two_rank_sharded_routing_equals_sequential_oracle():
    tmp = tempfile.mkdtemp()
    initfile, resfile = os.path.join(tmp, "init"), os.path.join(tmp, "res")
    mp.spawn(_worker, args=(initfile, resfile), nprocs=WORLD, join=True)
    res = [pickle.load(open(resfile + str(r), "rb")) for r in range(WORLD)]
    orc = coracle.TrieOracle()
    seq = 0
    for b, (avail, p0, p1) in enumerate(_shards()):
        buf, starts, lens, _ = traces.pack_strings(p0 + p1)
        oep, odepth, omatched = orc.route_batch(buf, starts, lens, avail, 77, seq)
        seq += len(p0) + len(p1)
        got_ep = res[0][0][b][0] + res[1][0][b][0]
        got_depth = res[0][0][b][1] + res[1][0][b][1]
        got_matched = res[0][0][b][2] + res[1][0][b][2]
        assert got_ep == oep.tolist(), b
        assert got_depth == odepth.tolist(), b
        assert got_matched == omatched.tolist(), b
    assert res[0][1] == res[1][1] == orc.dump()
