"""World-size-2 runs of the multi-GPU routing designs on CPUs (gloo).

Second test: the partitioned design as an algorithm (tests/fake_part_router.py).
First test:

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


def _part_worker(rank, initfile, resfile):
    import sys
    here = os.path.dirname(os.path.abspath(__file__))
    sys.path.insert(0, here)
    from fake_part_router import FakePartitionedRouter
    dist.init_process_group("gloo", init_method="file://" + initfile, rank=rank, world_size=WORLD)
    router = FakePartitionedRouter()
    results = []
    for avail, p0, p1 in _shards():
        results.append(router.route((p0, p1)[rank], avail, 77))
    with open(resfile + str(rank), "wb") as f:
        pickle.dump((results, router.dump()), f)
    dist.barrier()
    dist.destroy_process_group()


def test_two_rank_partitioned_routing_equals_sequential_oracle():
    """The partitioned design (bench.py --gpus N default; kernels in csrc/prefix.cu k_part_*)
    as an algorithm, in two real processes over gloo: ownership by first-chunk hash, each
    partition routed in global order, tie-break counter from the global position.  Decisions
    equal sequential routing of the rank-major batch and the union of the partitions is the
    sequential trie (root masks OR-ed, no node in two partitions)."""
    tmp = tempfile.mkdtemp()
    initfile, resfile = os.path.join(tmp, "init"), os.path.join(tmp, "res")
    mp.spawn(_part_worker, args=(initfile, resfile), nprocs=WORLD, join=True)
    res = [pickle.load(open(resfile + str(r), "rb")) for r in range(WORLD)]
    orc = coracle.TrieOracle()
    seq = 0
    for b, (avail, p0, p1) in enumerate(_shards()):
        buf, starts, lens, _ = traces.pack_strings(p0 + p1)
        oep, odepth, omatched = orc.route_batch(buf, starts, lens, avail, 77, seq)
        seq += len(p0) + len(p1)
        for k, want in enumerate((oep, odepth, omatched)):
            assert res[0][0][b][k] + res[1][0][b][k] == want.tolist(), (b, k)
    d0, d1 = res[0][1], res[1][1]
    assert not (set(d0) & set(d1)) - {()}
    merged = dict(d0)
    merged.update(d1)
    merged[()] = d0[()] | d1[()]
    assert merged == orc.dump()
    assert len(d0) > 1 and len(d1) > 1          # both partitions hold something
