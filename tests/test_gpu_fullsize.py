"""Parity at the sizes BASELINE.json quotes (configs C1, C3, C4-shaped), against the
C oracle (sequential pointer trie): decisions, match depths, candidate sets, and
size-independent properties of the final table."""
import numpy as np
import pytest
import torch

from b200_router import traces
from b200_router.packing import pack_arrays, pack_prompts
from oracle import coracle

from helpers import u64

pytestmark = pytest.mark.gpu


def test_c1_multi_round_qa_10k_prompts_4_endpoints():
    """configs[0]: 10K synthetic multi-round-QA prompts (1000 users x 10 rounds), 4 endpoints."""
    from b200_router.trie import GpuHashTrie
    prompts, _ = traces.multi_round_qa(users=1000, rounds=10, sys_len=1000, hist_len=200,
                                       answer_len=100)
    t = GpuHashTrie(cap_log2=21)
    orc = coracle.TrieOracle()
    seq = 0
    for r in range(10):                                   # one batch per round
        batch = prompts[r * 1000:(r + 1) * 1000]
        avail = 0xF if r not in (4, 5) else 0xD            # endpoint 1 away for two rounds
        buf, starts, lens, _ = traces.pack_strings(batch)
        ep, depth, matched = t.route(pack_arrays(buf, starts, lens, pin=False), avail, 1, seq)
        oep, odepth, omatched = orc.route_batch(buf, starts, lens, avail, 1, seq)
        seq += 1000
        assert (ep.cpu().numpy() == oep).all(), r
        assert (depth.cpu().numpy() == odepth).all(), r
        assert (u64(matched) == omatched).all(), r
    st = t.check_status(count_nodes=True)
    assert st["num_nodes"] == orc.num_nodes() - 1
    # multi-round property of the workload: a user's later rounds follow its first round
    assert t.dump() == orc.dump()
    t.close()


def test_c3_one_million_prompts_batch_64k():
    """configs[2] at full size: 1M prompts, 64 shared 2048-char system prompts, 8 endpoints,
    batches of 65536, endpoint 3 dropped for batches 5-6; through the batch pipeline."""
    from b200_router.trie import GpuHashTrie, RoutePipeline
    tr = traces.SharedPrefixTrace(n_total=1 << 20, batch=1 << 16, n_sys=64, sys_chars=2048,
                                  suffix_lo=256, suffix_hi=1024, seed=7)
    t = GpuHashTrie(cap_log2=24)
    pipe = RoutePipeline(t)
    orc = coracle.TrieOracle()
    seq = 0
    for b in range(tr.num_batches()):
        avail = 0xFF & ~(1 << 3) if b in (5, 6) else 0xFF
        buf, starts, lens = tr.batch_arrays(b)
        ep, depth, matched = pipe.route(pack_arrays(buf, starts, lens, pin=True), avail, 99, seq)
        oep, odepth, omatched = orc.route_batch(buf, starts, lens, avail, 99, seq)
        seq += len(starts)
        assert (ep.cpu().numpy() == oep).all(), b
        assert (depth.cpu().numpy() == odepth).all(), b
        assert (u64(matched) == omatched).all(), b
    st = t.check_status(count_nodes=True)
    assert st["num_nodes"] == orc.num_nodes() - 1
    # size-independent checks on the 5.6M-node table (no Python-level decode of it):
    arr, root = t.dump_raw()
    assert len(arr) == orc.num_nodes() - 1 and root == 0xFF
    _, oh, odepth_n, omask = orc.dump_arrays()
    # same multiset of (chunk hash, mask) over all nodes, and checksum of the keys is stable
    got = np.sort(arr[:, 2] * np.uint64(0x9E3779B97F4A7C15) ^ arr[:, 3])
    want = np.sort(oh[1:] * np.uint64(0x9E3779B97F4A7C15) ^ omask[1:])
    assert (got == want).all()
    # trie invariant child.mask subset-of parent.mask, checked through the parent links
    idx = arr[:, 0] & np.uint64(0xFFFFFFFF)
    order = np.argsort(idx)
    sidx, smask, spar = idx[order], arr[order, 3], arr[order, 1] & np.uint64(0xFFFFFFFF)
    assert (np.diff(sidx.astype(np.int64)) > 0).all()         # every node has its own log record
    is_top = spar == np.uint64(0xFFFFFFFF)
    pos = np.minimum(np.searchsorted(sidx, spar), len(sidx) - 1)
    assert ((sidx[pos] == spar) | is_top).all()               # prefix-closed: every parent exists
    assert (((smask & ~smask[pos]) == 0) | is_top).all()      # nested endpoint sets
    # runs: the lengths on the run heads add up to the node count
    heads = (arr[:, 1] >> np.uint64(63)) == 1
    assert int((arr[heads, 0] >> np.uint64(32)).sum()) == len(arr)
    t.close()


def test_c4_shape_long_context_64_endpoints():
    """configs[3] shape, scaled to one GPU: 32768-char prompts (256 chunks), 64 endpoints,
    groups of 64 prompts sharing half of their chunks."""
    from b200_router.trie import GpuHashTrie
    t = GpuHashTrie(cap_log2=21)
    orc = coracle.TrieOracle()
    seq = 0
    full = (1 << 64) - 1
    for b, (avail, frac) in enumerate([(full, 0.5), (full & ~0xFF00, 0.94), (full, 0.0)]):
        buf, starts, lens = traces.long_context_batch(n=1024, chars=32768, group=64,
                                                      shared_frac=frac, seed=b)
        ep, depth, matched = t.route(pack_arrays(buf, starts, lens, pin=False), avail, 5, seq)
        oep, odepth, omatched = orc.route_batch(buf, starts, lens, avail, 5, seq)
        seq += len(starts)
        assert (ep.cpu().numpy() == oep).all()
        assert (depth.cpu().numpy() == odepth).all()
        assert (u64(matched) == omatched).all()
    assert t.check_status(count_nodes=True)["num_nodes"] == orc.num_nodes() - 1
    t.close()
