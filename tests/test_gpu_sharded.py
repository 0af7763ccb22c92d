"""The multi-GPU paths on ONE device.  Replicated design (ShardedPrefixRouter over
b200r_xxh64_chunks + b200r_prefix_route_hashes): the ranks are threads of this process, each
with its own table replica, and the all-gathers are done through shared memory (the
orchestration code is the product's; only the collective is swapped).  Partitioned design
(b200r_part_*): the ranks run their phases in lock-step.  Results must equal routing the
rank-major concatenation sequentially (C oracle); all replicas must be identical.
The real NCCL path is exercised by `torchrun ... bench.py --gpus N`."""
import threading

import numpy as np
import pytest
import torch

from b200_router import traces
from b200_router.packing import pack_arrays, pack_prompts
from oracle import coracle

from helpers import u64

pytestmark = pytest.mark.gpu


class ThreadComm:
    def __init__(self, world, rank, shared):
        self.world, self.rank, self.sh = world, rank, shared

    def all_gather(self, out_flat, inp_flat):
        self.sh["slots"][self.rank] = inp_flat
        self.sh["barrier"].wait()
        out_flat.copy_(torch.cat([s.reshape(-1) for s in self.sh["slots"]]))
        torch.cuda.synchronize()
        self.sh["barrier"].wait()


def _run_ranks(world, batches, seed, cap_log2=17):
    """batches: [(avail, [per-rank packed prompts])] -> per-rank results and dumps."""
    from b200_router.distributed import ShardedPrefixRouter
    shared = {"barrier": threading.Barrier(world), "slots": [None] * world}
    results = [[] for _ in range(world)]
    dumps = [None] * world
    errors = []

    def work(rank):
        try:
            torch.cuda.set_device(0)
            r = ShardedPrefixRouter(cap_log2=cap_log2, device="cuda:0",
                                    comm=ThreadComm(world, rank, shared))
            seq = 0
            for avail, shards in batches:
                ep, depth, matched = r.route(shards[rank], avail, seed, seq)
                seq += sum(s.n for s in shards)
                results[rank].append((ep.cpu().numpy().copy(), depth.cpu().numpy().copy(),
                                      u64(matched).copy()))
            r.check_status()
            dumps[rank] = r.dump()
        except Exception as e:          # pragma: no cover
            errors.append(e)
            shared["barrier"].abort()

    ts = [threading.Thread(target=work, args=(k,)) for k in range(world)]
    [t.start() for t in ts]
    [t.join() for t in ts]
    if errors:
        raise errors[0]
    return results, dumps


def _check(world, batches_np, seed):
    """batches_np: [(avail, [(buf, starts, lens) per rank])]"""
    batches = [(a, [pack_arrays(*s, pin=False) for s in shards]) for a, shards in batches_np]
    results, dumps = _run_ranks(world, batches, seed)
    orc = coracle.TrieOracle()
    seq = 0
    for b, (avail, shards) in enumerate(batches_np):
        # sequential over the concatenation: feed the shards one after another
        off = 0
        for r, (buf, st, ln) in enumerate(shards):
            oep, odepth, omatched = orc.route_batch(buf, st, ln, avail, seed, seq + off)
            off += len(st)
            ep, depth, matched = results[r][b]
            assert (ep == oep).all(), (b, r)
            assert (depth == odepth).all(), (b, r)
            assert (matched == omatched).all(), (b, r)
        seq += off
    for d in dumps:
        assert d == orc.dump()


def test_sharded_heavy_sharing_two_and_three_ranks():
    for world in (2, 3):
        batches = []
        for avail_idx, prompts in traces.sharing_stress(n_batches=5, batch=90, n_endpoints=7,
                                                        max_depth=6, alphabet=2, seed=40 + world):
            cuts = np.linspace(0, len(prompts), world + 1).astype(int)
            shards = []
            for r in range(world):
                buf, st, ln, _ = traces.pack_strings(prompts[cuts[r]:cuts[r + 1]])
                shards.append((buf, st, ln))
            batches.append((sum(1 << i for i in avail_idx), shards))
        _check(world, batches, seed=5)


def test_sharded_shared_prefix_trace_with_churn():
    world = 4
    tr = traces.SharedPrefixTrace(n_total=16000, batch=1000, n_sys=8, sys_chars=512,
                                  suffix_lo=100, suffix_hi=300, seed=9)
    batches = []
    for g in range(4):
        avail = 0xFF & ~(1 << 1) if g in (1, 2) else 0xFF
        batches.append((avail, [tr.batch_arrays(g * world + r) for r in range(world)]))
    _check(world, batches, seed=11)


def test_sharded_uneven_and_empty_shards():
    world = 2
    _, prompts = traces.sharing_stress(n_batches=1, batch=50, seed=8)[0]
    shards0 = [traces.pack_strings(prompts[:50])[:3], traces.pack_strings([])[:3]]
    shards1 = [traces.pack_strings(prompts[:7])[:3], traces.pack_strings(prompts[7:])[:3]]
    _check(world, [(0x3F, shards0), (0x15, shards1)], seed=2)


def _merge_partitions(dumps):
    out = {}
    for d in dumps:
        for path, m in d.items():
            if path == ():
                out[path] = out.get(path, 0) | m
            else:
                assert path not in out, "a node lives in two partitions"
                out[path] = m
    return out


def _check_partitioned_lockstep(world, batches_np, seed):
    """Partitioned-table path on ONE device, ranks in lock-step (nothing ever spins)."""
    from b200_router.distributed import PartitionedPrefixRouter as R
    max_req = max(len(s[1]) for _, shards in batches_np for s in shards)
    batches = [(a, [pack_arrays(*s, pin=False) for s in shards]) for a, shards in batches_np]
    max_chunks = max(p.max_chunks for _, shards in batches for p in shards)
    nbytes = R.symm_bytes_for(world, max_req, max_chunks)
    bufs = [torch.zeros(nbytes, dtype=torch.uint8, device="cuda:0") for _ in range(world)]
    ranks = [R(cap_log2=17, device="cuda:0", max_requests=max_req, max_chunks=max_chunks,
               _sim={"world": world, "rank": r, "buffers": bufs}) for r in range(world)]
    orc = coracle.TrieOracle()
    seq = 0
    owned_total = np.zeros(world, dtype=np.int64)
    for avail, shards in batches:
        for r in range(world):
            ranks[r].phase_send(shards[r])
        for r in range(world):
            ranks[r].phase_route(avail, seed)
        outs = [ranks[r].phase_finish() for r in range(world)]
        torch.cuda.synchronize()
        off = 0
        for r, p in enumerate(shards):
            buf, st, ln = p.bytes.cpu().numpy(), p.starts.cpu().numpy(), p.lens.cpu().numpy()
            oep, odepth, omatched = orc.route_batch(buf, st, ln, avail, seed, seq + off)
            off += p.n
            ep, depth, matched = outs[r]
            assert (ep.cpu().numpy() == oep).all(), r
            assert (depth.cpu().numpy() == odepth).all(), r
            assert (u64(matched) == omatched).all(), r
        seq += off
    for rk in ranks:
        rk.check_status()
    assert _merge_partitions([rk.dump() for rk in ranks]) == orc.dump()
    sizes = [len(rk.dump()) for rk in ranks]
    assert sum(1 for s in sizes if s > 1) >= min(2, world)      # really spread over partitions


def test_partitioned_lockstep_heavy_sharing():
    for world in (2, 3):
        batches = []
        for avail_idx, prompts in traces.sharing_stress(n_batches=5, batch=90, n_endpoints=7,
                                                        max_depth=6, alphabet=3, seed=90 + world):
            cuts = np.linspace(0, len(prompts), world + 1).astype(int)
            shards = [traces.pack_strings(prompts[cuts[r]:cuts[r + 1]])[:3] for r in range(world)]
            batches.append((sum(1 << i for i in avail_idx), shards))
        _check_partitioned_lockstep(world, batches, seed=8)


def test_partitioned_lockstep_shared_prefix_and_empty_shard():
    world = 4
    tr = traces.SharedPrefixTrace(n_total=16000, batch=1000, n_sys=16, sys_chars=512,
                                  suffix_lo=100, suffix_hi=300, seed=23)
    batches = []
    for g in range(4):
        avail = 0xFF & ~(1 << 1) if g in (1, 2) else 0xFF
        shards = [tr.batch_arrays(g * world + r) for r in range(world)]
        if g == 2:
            shards[1] = traces.pack_strings([])[:3]
        batches.append((avail, shards))
    _check_partitioned_lockstep(world, batches, seed=14)


def test_partitioned_lockstep_with_the_staged_send_path():
    """ASCII batches normally have their hashes stored straight into the owners' buffers by the
    hashing kernel; b200r_set_part_staged(1) forces the compaction + shipment kernels that
    non-ASCII batches take.  Same decisions and tables either way."""
    from b200_router import _lib
    _lib.check(_lib.lib().b200r_set_part_staged(1))
    try:
        test_partitioned_lockstep_shared_prefix_and_empty_shard()
        test_partitioned_lockstep_heavy_sharing()
    finally:
        _lib.check(_lib.lib().b200r_set_part_staged(0))


def test_partitioned_lockstep_utf8_and_unaligned_prompts():
    """Non-ASCII batches take the staged send path (chunk boundaries come from the code-point
    scan); ASCII prompts packed without alignment have their first chunk hashed by the generic
    path of the placing prep kernel.  Both against the oracle, next to ordinary batches."""
    import random
    world = 2

    def tight_arrays(ps):                     # CSR-tight: unaligned starts
        blobs = [p.encode("ascii") for p in ps]
        lens = np.array([len(b) for b in blobs], dtype=np.int32)
        starts = np.zeros(len(blobs), dtype=np.int64)
        starts[1:] = np.cumsum(lens[:-1])
        return np.frombuffer(b"".join(blobs) + b"\0" * 16, dtype=np.uint8).copy(), starts, lens

    rng = random.Random(5)
    heads = ["systémé prompt " * 20, "Ünïcödé ✓ header " * 17, "plain ascii header " * 15]

    def prompts(k, ascii_only):
        out = []
        for _ in range(k):
            h = heads[2] if ascii_only else rng.choice(heads)
            tail = "".join(rng.choice("abcdé✓xyz" if not ascii_only else "abcdxyz") for _ in range(rng.randint(0, 400)))
            out.append(h + tail)
        return out

    batches = []
    for g in range(5):
        shards = []
        for r in range(world):
            if g % 2 == 0:
                ps = prompts(150, ascii_only=False)
                shards.append(traces.pack_strings(ps))                  # (buf, starts, lens, nchars)
            else:
                ps = prompts(150, ascii_only=True)
                shards.append(tight_arrays(ps) if r == 0 else traces.pack_strings(ps)[:3])
        batches.append((0x3F if g != 2 else 0x2F, shards))
    _check_partitioned_lockstep(world, batches, seed=21)
