"""Parity of the CUDA prefix-routing path (through the C ABI) with the oracle
and with the golden fixtures generated from the unmodified reference."""
import asyncio

import numpy as np
import pytest
import torch

from oracle import coracle, prefix_ref
from b200_router import traces
from b200_router.packing import pack_arrays, pack_prompts

from helpers import bits, tight_pack, trie_as_sorted_list, u64

pytestmark = pytest.mark.gpu


@pytest.fixture()
def trie():
    from b200_router.trie import GpuHashTrie
    t = GpuHashTrie(cap_log2=18)
    yield t
    t.close()


def _oracle_hashes(prompts):
    return [coracle.xxh64_chunks(p.encode("utf-8"), 128) for p in prompts]


def _check_hashes(trie, packed, prompts):
    base, hs = trie.hash_chunks(packed)
    base = base.cpu().numpy()
    hs = u64(hs)
    want = _oracle_hashes(prompts)
    assert base[0] == 0
    for i, w in enumerate(want):
        assert base[i + 1] - base[i] == len(w), (i, len(prompts[i]))
        assert (hs[base[i]:base[i + 1]] == w).all(), i


def test_xxh64_known_answers(trie):
    kat = {"": None, "a": 0xd24ec4f1a98c6e5b, "abc": 0x44bc2cf5ad770999,
           "1" * 128: 0x2ceab0ab9ef63228, "a" * 128: 0x6550288941193272,
           "hi " * 42 + "hi": 0x380de794396a43ea, "é" * 128: 0x9b2210ff98cf44b0}
    prompts = list(kat)
    base, hs = trie.hash_chunks(pack_prompts(prompts))
    base, hs = base.cpu().numpy(), u64(hs)
    for i, p in enumerate(prompts):
        if kat[p] is None:
            assert base[i + 1] == base[i]
        else:
            assert base[i + 1] - base[i] == 1 and int(hs[base[i]]) == kat[p], p


def test_xxh64_golden_chunking(trie, golden_xxh):
    texts = [c["text"] for c in golden_xxh["chunked"]]
    base, hs = trie.hash_chunks(pack_prompts(texts))
    base, hs = base.cpu().numpy(), u64(hs)
    for i, c in enumerate(golden_xxh["chunked"]):
        assert ["%016x" % int(h) for h in hs[base[i]:base[i + 1]]] == c["hashes"], i


def test_xxh64_all_lengths_ascii_aligned(trie):
    rng = np.random.default_rng(0)
    prompts = ["".join(map(chr, rng.integers(32, 127, size=n))) for n in range(0, 520)]
    prompts += ["".join(map(chr, rng.integers(32, 127, size=n))) for n in (1000, 4096, 4097, 33000)]
    _check_hashes(trie, pack_prompts(prompts), prompts)


def test_xxh64_unaligned_and_utf8(trie):
    rng = np.random.default_rng(1)
    alphabet = list("abcXYZ 01") + ["é", "ü", "日", "本", "🙂", "߿", "ࠀ", "\U00010000"]
    prompts = ["".join(rng.choice(alphabet, size=n)) for n in list(range(0, 300, 7)) + [128, 129, 1000]]
    prompts += ["", "x" * 5, ""]
    _check_hashes(trie, pack_prompts(prompts), prompts)        # aligned, mixed ascii / utf-8
    _check_hashes(trie, tight_pack(prompts), prompts)          # CSR-tight: unaligned starts


def test_xxh64_device_counts_chars_when_nchars_missing(trie):
    prompts = ["é" * 300, "abc", "日本語" * 100, ""]
    p = pack_prompts(prompts)
    p.nchars = None          # all_ascii stays False -> library must count code points itself
    _check_hashes(trie, p, prompts)


def _replay_golden(trie, tr):
    seq = 0
    dec, looks = [], []
    for b in tr["batches"]:
        if tr["json_inputs"] is not None:
            prompts = [prefix_ref.flatten_prompt(tr["json_inputs"][i]) for i in b["prompts"]]
        else:
            prompts = b["prompts"]
        avail = sum(1 << i for i in b["avail"])
        ep, depth, matched = trie.route(pack_prompts(prompts), avail, tr["seed"], seq)
        seq += len(prompts)
        dec += ep.cpu().tolist()
        looks += [[int(d), bits(m)] for d, m in zip(depth.cpu().tolist(), u64(matched))]
    return dec, looks


def test_route_golden_traces(golden_prefix):
    """Decisions, per-request lookup results and final trie state equal the
    unmodified reference's on every golden trace (batched per availability set)."""
    from b200_router.trie import GpuHashTrie
    for tr in golden_prefix["traces"]:
        t = GpuHashTrie(cap_log2=14)
        dec, looks = _replay_golden(t, tr)
        assert dec == tr["decisions"], tr["name"]
        if tr["lookups"] is not None:
            assert looks == tr["lookups"], tr["name"]
        st = t.check_status(count_nodes=True)
        assert trie_as_sorted_list(t.dump()) == tr["trie"], tr["name"]
        assert st["num_nodes"] == len(tr["trie"]) - 1
        t.close()


def test_route_golden_traces_one_request_per_batch(golden_prefix):
    """Same traces with batch size 1: the degenerate sequential schedule."""
    from b200_router.trie import GpuHashTrie
    tr = golden_prefix["traces"][1]
    t = GpuHashTrie(cap_log2=14)
    seq, dec = 0, []
    for b in tr["batches"]:
        avail = sum(1 << i for i in b["avail"])
        for p in b["prompts"]:
            ep, _, _ = t.route(pack_prompts([p]), avail, tr["seed"], seq)
            seq += 1
            dec.append(int(ep.cpu()[0]))
    assert dec == tr["decisions"]
    assert trie_as_sorted_list(t.dump()) == tr["trie"]
    t.close()


def _route_vs_oracle(trie, batches, seed, chunk=128):
    """batches: [(avail_mask, buf, starts, lens)] -> compares everything."""
    orc = coracle.TrieOracle(chunk)
    seq = 0
    for avail, buf, starts, lens in batches:
        ep, depth, matched = trie.route(pack_arrays(buf, starts, lens, pin=False), avail, seed, seq)
        oep, odepth, omatched = orc.route_batch(buf, starts, lens, avail, seed, seq)
        seq += len(starts)
        assert (ep.cpu().numpy() == oep).all()
        assert (depth.cpu().numpy() == odepth).all()
        assert (u64(matched) == omatched).all()
    trie.check_status()
    assert trie.dump() == orc.dump()
    return orc


def test_route_shared_prefix_with_churn_vs_oracle():
    """C3-shaped trace, scaled down: shared system prompts + unique suffixes,
    one endpoint dropped for two batches and restored (multi-bit masks)."""
    from b200_router.trie import GpuHashTrie
    tr = traces.SharedPrefixTrace(n_total=24000, batch=4096, n_sys=16, sys_chars=512,
                                  suffix_lo=100, suffix_hi=400, seed=3)
    full = 0xFF
    batches = []
    for b in range(tr.num_batches()):
        avail = full & ~(1 << 2) if b in (2, 3) else full
        batches.append((avail,) + tr.batch_arrays(b))
    t = GpuHashTrie(cap_log2=19)
    _route_vs_oracle(t, batches, seed=42)
    t.close()


def test_route_heavy_sharing_random_vs_oracle():
    """Many small random traces with tiny alphabets: long leader chains, every
    availability subset, empty prompts, partial last chunks."""
    from b200_router.trie import GpuHashTrie
    for s in range(12):
        t = GpuHashTrie(cap_log2=12)
        batches = []
        for avail_idx, prompts in traces.sharing_stress(n_batches=5, batch=150, n_endpoints=7,
                                                        max_depth=6, alphabet=2, seed=1000 + s):
            buf, starts, lens, _ = traces.pack_strings(prompts)
            batches.append((sum(1 << i for i in avail_idx), buf, starts, lens))
        _route_vs_oracle(t, batches, seed=s)
        t.close()


def test_route_pipeline_overlapped_batches_vs_oracle():
    """RoutePipeline (hash stage of batch b+1 overlapping the route stage of batch b,
    host inputs copied on the hash stream) gives the sequential results."""
    from b200_router.trie import GpuHashTrie, RoutePipeline
    tr = traces.SharedPrefixTrace(n_total=60000, batch=6000, n_sys=16, sys_chars=512,
                                  suffix_lo=100, suffix_hi=400, seed=21)
    for host_inputs in (False, True):
        t = GpuHashTrie(cap_log2=19)
        pipe = RoutePipeline(t)
        orc = coracle.TrieOracle()
        seq, pending = 0, []
        for b in range(tr.num_batches()):
            avail = 0xFF & ~(1 << 5) if b in (3, 4, 5) else 0xFF
            buf, starts, lens = tr.batch_arrays(b)
            if b % 3 == 1:            # small batches take the pipeline's single-stream path
                m = 40 if b == 1 else 900
                starts, lens = starts[:m], lens[:m]
            p = pack_arrays(buf, starts, lens, pin=host_inputs)
            if not host_inputs:
                p = p.to(t.device)
                torch.cuda.synchronize()
            ep, depth, matched = pipe.route(p, avail, 3, seq, inputs_ready=not host_inputs)
            want = orc.route_batch(buf, starts, lens, avail, 3, seq)
            seq += len(starts)
            pending.append((ep.clone(), depth.clone(), matched.clone(), want))
        torch.cuda.synchronize()
        for ep, depth, matched, (oep, odepth, omatched) in pending:
            assert (ep.cpu().numpy() == oep).all()
            assert (depth.cpu().numpy() == odepth).all()
            assert (u64(matched) == omatched).all()
        t.check_status()
        assert t.dump() == orc.dump()
        t.close()


def test_long_and_short_prompts_mixed_vs_oracle():
    """Prompts of 45 chunks (several 32-depth windows of the walk, runs longer than a warp) next
    to C3-shaped ones with suffixes down to one character, with availability churn."""
    from b200_router.trie import GpuHashTrie
    bufL, stL, lnL = traces.long_context_batch(n=300, chars=128 * 45 + 77, group=10,
                                               shared_frac=0.5, seed=15)
    tr = traces.SharedPrefixTrace(n_total=9000, batch=3000, n_sys=16, sys_chars=512,
                                  suffix_lo=1, suffix_hi=400, seed=22)
    batches = [(0xFF, bufL, stL, lnL)] + [((0xFB if b == 1 else 0xFF),) + tr.batch_arrays(b)
                                          for b in range(3)]
    t = GpuHashTrie(cap_log2=18)
    _route_vs_oracle(t, batches, seed=4)
    t.close()


def test_route_from_hashes_and_on_a_restored_table_vs_oracle():
    """b200r_prefix_route_hashes (route stage from caller-held chunk hashes) equals routing the
    prompts; a table restored from a dump keeps routing like the original (same layout), and a
    dump can only be restored into an empty table."""
    from b200_router import B200RouterError
    from b200_router.trie import GpuHashTrie
    tr = traces.SharedPrefixTrace(n_total=8000, batch=2000, n_sys=8, sys_chars=1024,
                                  suffix_lo=1, suffix_hi=600, seed=31)
    bufL, stL, lnL = traces.long_context_batch(n=64, chars=128 * 40 + 9, group=8, shared_frac=0.7, seed=2)
    orc = coracle.TrieOracle()
    t = GpuHashTrie(cap_log2=17)
    seq = 0
    first = [(0xFF, bufL, stL, lnL)] + [(0xFF,) + tr.batch_arrays(b) for b in range(2)]
    for k, (avail, buf, starts, lens) in enumerate(first):
        p = pack_arrays(buf, starts, lens, pin=False)
        if k == 1:                                    # this batch through the hashes entry point
            base, hs = t.hash_chunks(p)
            ep, depth, matched = t.route_hashes(hs, base, p.n, p.max_chunks, avail, 6, seq)
        else:
            ep, depth, matched = t.route(p, avail, 6, seq)
        oep, odepth, omatched = orc.route_batch(buf, starts, lens, avail, 6, seq)
        seq += len(starts)
        assert (ep.cpu().numpy() == oep).all(), k
        assert (depth.cpu().numpy() == odepth).all(), k
        assert (u64(matched) == omatched).all(), k
    arr, root = t.dump_raw()
    t2 = GpuHashTrie(cap_log2=17)
    t2.restore_entries(torch.from_numpy(arr.view(np.int64).copy()), root)
    assert t2.dump() == orc.dump()
    assert t2.status()["num_nodes"] == t.status()["num_nodes"]
    for avail, buf, starts, lens in [(0xF7,) + tr.batch_arrays(2), (0xFF, bufL, stL, lnL),
                                     (0xFF,) + tr.batch_arrays(3)]:
        ep, depth, matched = t2.route(pack_arrays(buf, starts, lens, pin=False), avail, 6, seq)
        oep, odepth, omatched = orc.route_batch(buf, starts, lens, avail, 6, seq)
        seq += len(starts)
        assert (ep.cpu().numpy() == oep).all()
        assert (depth.cpu().numpy() == odepth).all()
        assert (u64(matched) == omatched).all()
    t2.check_status()
    assert t2.dump() == orc.dump()
    with pytest.raises(B200RouterError):
        t2.restore_entries(torch.from_numpy(arr.view(np.int64).copy()), root)
    t.close()
    t2.close()


def test_latency_path_one_request_per_launch(golden_prefix):
    """b200r_prefix_route_one (one launch per request, host bytes in, host result out) equals
    the unmodified reference on every golden trace -- decisions, lookup results, final trie;
    prompts it does not take (non-ASCII) go through the batch path in between."""
    from b200_router.trie import GpuHashTrie
    taken = 0
    for tr in golden_prefix["traces"]:
        t = GpuHashTrie(cap_log2=14)
        seq, dec, looks = 0, [], []
        for b in tr["batches"]:
            avail = sum(1 << i for i in b["avail"])
            for item in b["prompts"]:
                p = prefix_ref.flatten_prompt(tr["json_inputs"][item]) if tr["json_inputs"] is not None else item
                r = t.route_one(p.encode("utf-8"), avail, tr["seed"], seq)
                if r is None:
                    assert not p.isascii(), tr["name"]
                    ep, depth, matched = t.route(pack_prompts([p]), avail, tr["seed"], seq)
                    r = (int(ep.cpu()[0]), int(depth.cpu()[0]), int(u64(matched)[0]), 0)
                else:
                    assert p.isascii() and r[3] == 0
                    taken += 1
                seq += 1
                dec.append(r[0])
                looks.append([r[1], bits(r[2])])
        assert dec == tr["decisions"], tr["name"]
        if tr["lookups"] is not None:
            assert looks == tr["lookups"], tr["name"]
        assert trie_as_sorted_list(t.dump()) == tr["trie"], tr["name"]
        t.close()
    assert taken > 100


def test_latency_path_between_batches_vs_oracle():
    """Lone requests (latency path) interleaved with batches on one table, against the C
    oracle: long prompts (several staging passes), an empty prompt, availability churn."""
    from b200_router.trie import GpuHashTrie
    tr = traces.SharedPrefixTrace(n_total=3000, batch=1000, n_sys=6, sys_chars=1024,
                                  suffix_lo=1, suffix_hi=500, seed=77)
    bufL, stL, lnL = traces.long_context_batch(n=12, chars=128 * 300 + 5, group=4, shared_frac=0.6, seed=5)
    orc = coracle.TrieOracle()
    t = GpuHashTrie(cap_log2=17)
    seq = 0

    def one(buf, start, ln, avail):
        nonlocal seq
        raw = bytes(buf[start:start + ln])
        r = t.route_one(raw, avail, 9, seq)
        assert r is not None
        oep, odepth, omatched = orc.route_batch(np.frombuffer(raw, dtype=np.uint8) if ln else np.zeros(16, np.uint8),
                                                np.zeros(1, np.int64), np.array([ln], np.int32), avail, 9, seq)
        seq += 1
        assert (r[0], r[1], r[2]) == (int(oep[0]), int(odepth[0]), int(omatched[0]))

    def batch(buf, starts, lens, avail):
        nonlocal seq
        ep, depth, matched = t.route(pack_arrays(buf, starts, lens, pin=False), avail, 9, seq)
        oep, odepth, omatched = orc.route_batch(buf, starts, lens, avail, 9, seq)
        seq += len(starts)
        assert (ep.cpu().numpy() == oep).all() and (depth.cpu().numpy() == odepth).all()
        assert (u64(matched) == omatched).all()

    b0, b1, b2 = (tr.batch_arrays(k) for k in range(3))
    for i in range(40):                                # cold table: lone requests create the runs
        one(b0[0], int(b0[1][i]), int(b0[2][i]), 0xFF)
    batch(*b1, 0xFF)
    for i in range(12):                                # 301 chunks each: two staging passes
        one(bufL, int(stL[i]), int(lnL[i]), 0xFF if i % 3 else 0x0F)
    one(b0[0], 0, 0, 0xFF)                             # empty prompt
    for i in range(40, 200):
        one(b0[0], int(b0[1][i]), int(b0[2][i]), 0xF7 if i % 5 == 0 else 0xFF)
    batch(*b2, 0xFF)
    batch(bufL, stL, lnL, 0xFF)
    t.check_status()
    assert t.dump() == orc.dump()
    t.close()


def test_batch_epoch_wraps_without_losing_the_table():
    """The per-node first-toucher words are tagged with a 32-bit batch epoch.  When it runs out
    the library sweeps those words and restarts the count; learned prefixes stay
    (routing_logic.py:352: the reference trie has no lifetime limit)."""
    from b200_router import _lib
    from b200_router.trie import GpuHashTrie
    t = GpuHashTrie(cap_log2=12)
    batches = []
    for avail_idx, prompts in traces.sharing_stress(n_batches=6, batch=150, n_endpoints=7,
                                                    max_depth=6, alphabet=2, seed=4242):
        buf, starts, lens, _ = traces.pack_strings(prompts)
        batches.append((sum(1 << i for i in avail_idx), buf, starts, lens))
    orc = coracle.TrieOracle()
    seq = 0
    for b, (avail, buf, starts, lens) in enumerate(batches):
        if b == 2:
            _lib.check(_lib.lib().b200r_trie_set_epoch(t._h, 0xFFFFFFEF))    # one batch before the limit
        ep, depth, matched = t.route(pack_arrays(buf, starts, lens, pin=False), avail, 11, seq)
        oep, odepth, omatched = orc.route_batch(buf, starts, lens, avail, 11, seq)
        seq += len(starts)
        assert (ep.cpu().numpy() == oep).all(), b
        assert (depth.cpu().numpy() == odepth).all(), b
        assert (u64(matched) == omatched).all(), b
    t.check_status()
    assert t.dump() == orc.dump()
    t.close()


def test_retire_endpoints_clears_their_bits_everywhere():
    from b200_router.trie import GpuHashTrie
    t = GpuHashTrie(cap_log2=14)
    _, prompts = traces.sharing_stress(n_batches=1, batch=400, n_endpoints=6, seed=5)[0]
    p = pack_prompts(prompts)
    t.route(p, 0x3F, 2, 0)
    before = t.dump()
    assert any(m & 0b1010 for m in before.values())
    t.retire_endpoints(0b1010)
    after = t.dump()
    assert set(after) == set(before)
    assert all(after[k] == before[k] & ~0b1010 for k in before)
    t.close()


def test_route_long_prompts_many_blocks_vs_oracle():
    """Prompts of more than 32 chunks (several windows of the warp walker, runs spanning windows)."""
    from b200_router.trie import GpuHashTrie
    buf, starts, lens = traces.long_context_batch(n=96, chars=128 * 70 + 5, group=8,
                                                  shared_frac=0.6, seed=5)
    t = GpuHashTrie(cap_log2=15)
    half = len(starts) // 2
    _route_vs_oracle(t, [((1 << 64) - 1, buf, starts[:half], lens[:half]),
                         (0x0F0F, buf, starts[half:], lens[half:]),
                         ((1 << 64) - 1, buf, starts, lens)], seed=9)
    t.close()


def test_lookup_and_insert_batches_vs_oracle(trie):
    """a2 / a3 entry points on their own: externally chosen endpoints."""
    rng = np.random.default_rng(7)
    orc = coracle.TrieOracle()
    for rnd in range(4):
        _, prompts = traces.sharing_stress(n_batches=1, batch=300, max_depth=5, alphabet=3,
                                           seed=50 + rnd)[0]
        buf, starts, lens, _ = traces.pack_strings(prompts)
        p = pack_arrays(buf, starts, lens, pin=False)
        ep = rng.integers(0, 9, size=len(prompts)).astype(np.int32)
        trie.insert(p, torch.from_numpy(ep))
        orc.insert_batch(buf, starts, lens, ep)
        for avail in (0x1FF, 0x5, 0x100, 0x0):
            d, m = trie.longest_prefix_match(p, avail)
            od, om = orc.lookup_batch(buf, starts, lens, avail)
            assert (d.cpu().numpy() == od).all()
            assert (u64(m) == om).all()
    assert trie.dump() == orc.dump()


def test_dump_restore_roundtrip(trie):
    from b200_router.trie import GpuHashTrie
    _, prompts = traces.sharing_stress(n_batches=1, batch=200, seed=77)[0]
    p = pack_prompts(prompts)
    trie.route(p, 0x3F, 1, 0)
    arr, root = trie.dump_raw()
    t2 = GpuHashTrie(cap_log2=13)
    t2.restore_entries(torch.from_numpy(arr.view(np.int64).copy()), root)
    assert t2.dump() == trie.dump()
    assert t2.status()["num_nodes"] == trie.status()["num_nodes"]
    d1, m1 = trie.longest_prefix_match(p, 0x3F)
    d2, m2 = t2.longest_prefix_match(p, 0x3F)
    assert torch.equal(d1, d2) and torch.equal(m1, m2)
    t2.close()


def test_table_full_is_reported():
    from b200_router import B200RouterError
    from b200_router.trie import GpuHashTrie
    t = GpuHashTrie(cap_log2=4)          # 16 entries
    prompts = ["%03d" % i * 100 for i in range(40)]
    t.route(pack_prompts(prompts), 0x3, 0, 0)
    with pytest.raises(B200RouterError):
        t.check_status()
    t.close()


def test_router_e2e_prefix_families():
    """The reference's e2e assertion (tests/e2e/test-routing.py:299-418): every
    nested-prompt family lands on one endpoint."""
    from b200_router import routers
    routers._reset_singletons_for_tests()
    r = routers.initialize_routing_logic("prefixaware")
    assert routers.get_routing_logic() is r and isinstance(r, routers.PrefixAwareRouter)

    class Ep:
        def __init__(self, url): self.url = url
    eps = [Ep("http://localhost:8001"), Ep("http://localhost:8002")]
    c = 128
    fams = [["1" * c, "1" * c + "2" * c, "1" * c + "2" * c + "3" * c],
            ["2" * c + "3" * c + "4" * c, "2" * c + "3" * c, "2" * c],
            ["5" * c, "5" * c + "6" * c, "5" * c + "6" * c + "7" * c]]

    async def go():
        seen = []
        for fam in fams:
            urls = set()
            for sample in fam:
                for _ in range(5):
                    urls.add(await r.route_request(eps, None, None, None,
                                                   {"messages": [{"role": "user", "content": sample}]}))
            seen.append(urls)
        return seen
    for urls in asyncio.run(go()):
        assert len(urls) == 1
    with pytest.raises(KeyError):
        asyncio.run(r.route_request(eps, None, None, None, {"nothing": 1}))
    routers._reset_singletons_for_tests()


def test_checkpoint_save_load(tmp_path):
    from b200_router.trie import GpuHashTrie
    _, prompts = traces.sharing_stress(n_batches=1, batch=300, max_depth=6, seed=123)[0]
    t = GpuHashTrie(cap_log2=13)
    p = pack_prompts(prompts)
    t.route(p, 0x3F, 9, 0)
    path = str(tmp_path / "trie.npz")
    t.save(path)
    t2 = GpuHashTrie.load(path)
    assert t2.dump() == t.dump()
    # both continue identically
    a = t.route(p, 0x15, 9, 300)
    b = t2.route(p, 0x15, 9, 300)
    assert all(torch.equal(x, y) for x, y in zip(a, b))
    assert t2.dump() == t.dump()
    bare = str(tmp_path / "checkpoint_without_suffix")          # np.savez would append ".npz"
    t.save(bare)
    t3 = GpuHashTrie.load(bare)
    assert t3.dump() == t.dump()
    t.close(); t2.close(); t3.close()


def test_router_checkpoint_keeps_registry_and_counter(tmp_path):
    """PrefixAwareRouter.save/load: the urls behind the mask bits, the chunking mode and the
    tie-break counter travel with the table, so a reloaded router continues like the original."""
    from b200_router import routers

    class Ep:
        def __init__(self, url):
            self.url = url
    eps = [Ep(u) for u in traces.endpoint_urls(5)]
    _, prompts = traces.sharing_stress(n_batches=1, batch=200, max_depth=5, seed=5)[0]
    routers._reset_singletons_for_tests()
    r = routers.PrefixAwareRouter(cap_log2=13, seed=3)
    first = r.route_batch(eps[::-1], prompts[:100])             # bits assigned in reverse url order
    path = str(tmp_path / "router_state")
    r.save(path)
    want = r.route_batch(eps, prompts[100:])
    one = r.route_one(eps, prompts[0])
    routers._reset_singletons_for_tests()
    r2 = routers.PrefixAwareRouter(cap_log2=13, seed=99)
    r2.load(path)
    assert r2.route_batch(eps, prompts[100:]) == want
    assert r2.route_one(eps, prompts[0]) == one
    assert set(first) <= {e.url for e in eps}
    routers._reset_singletons_for_tests()


def test_router_micro_batched_concurrent_requests_equal_reference(golden_prefix):
    """f1: concurrent route_request coroutines are micro-batched; the outcome equals the
    reference's one-by-one routing of the same arrival order (golden decisions)."""
    from b200_router import routers
    tr = [t for t in golden_prefix["traces"] if t["name"] == "e2e_families"][0]
    routers._reset_singletons_for_tests()
    r = routers.PrefixAwareRouter(cap_log2=14, seed=tr["seed"], micro_batch_ms=20, micro_batch_max=64)

    class Ep:
        def __init__(self, url): self.url = url
    eps = [Ep(u) for u in tr["urls"]]
    prompts = tr["batches"][0]["prompts"]

    async def go():
        tasks = [asyncio.ensure_future(r.route_request(eps, None, None, None, {"prompt": p}))
                 for p in prompts]
        return await asyncio.gather(*tasks)
    urls = asyncio.run(go())
    assert [tr["urls"].index(u) for u in urls] == tr["decisions"]
    assert 1 < r.batcher.batches_routed <= (len(prompts) + 63) // 64 + 1
    routers._reset_singletons_for_tests()


def test_every_claim_decomposition_agrees_with_the_oracle():
    """The claim step can walk a request with one thread, 8, 16 or 32 lanes
    (b200r_set_claim_mode); every choice must equal the oracle -- heavy sharing inside batches
    (creation races on the same child, long leader chains), prompts longer than a tile and
    longer than a warp, runs that other requests leave in the middle."""
    from b200_router import _lib
    from b200_router.trie import GpuHashTrie
    bufL, stL, lnL = traces.long_context_batch(n=200, chars=128 * 70 + 3, group=8, shared_frac=0.4, seed=33)
    tr = traces.SharedPrefixTrace(n_total=12000, batch=4000, n_sys=8, sys_chars=1024,
                                  suffix_lo=1, suffix_hi=500, seed=77)
    try:
        for mode in (1, 2, 3, 4):
            _lib.check(_lib.lib().b200r_set_claim_mode(mode))
            for s in range(3):
                t = GpuHashTrie(cap_log2=17)
                batches = [((1 << 64) - 1, bufL, stL, lnL)] if s == 0 else []
                if s == 1:
                    batches += [((0xFB if b == 1 else 0xFF),) + tr.batch_arrays(b) for b in range(3)]
                for avail_idx, prompts in traces.sharing_stress(n_batches=4, batch=200, n_endpoints=7,
                                                                max_depth=6, alphabet=2, seed=300 + s):
                    buf, starts, lens, _ = traces.pack_strings(prompts)
                    batches.append((sum(1 << i for i in avail_idx), buf, starts, lens))
                _route_vs_oracle(t, batches, seed=s)
                t.close()
    finally:
        _lib.check(_lib.lib().b200r_set_claim_mode(0))


def test_byte_chunk_mode_reproduces_the_gateway_picker(trie, golden_go):
    """chunk_by="bytes" (+ from_root_endpoints for lookups) against the restated gateway picker
    (tests/golden/go_picker_kat.json; prefix_aware_picker.go:154-213): chunk hashes of UTF-8
    text cut at byte 128 -- inside multi-byte sequences -- and the insert / match trace."""
    texts = [bytes.fromhex(c["utf8_hex"]).decode() for c in golden_go["chunks"]]
    base, hs = trie.hash_chunks(pack_prompts(texts, chunk_by="bytes"))
    base, hs = base.cpu().numpy(), u64(hs)
    for i, c in enumerate(golden_go["chunks"]):
        assert ["%016x" % int(h) for h in hs[base[i]:base[i + 1]]] == c["hashes"], i
    eps = golden_go["endpoints"]
    for op in golden_go["ops"]:
        p = pack_prompts([op["key"]], chunk_by="bytes")
        if op["op"] == "insert":
            trie.insert(p, torch.tensor([eps.index(op["endpoint"])], dtype=torch.int32))
        else:
            avail = sum(1 << eps.index(e) for e in op["avail"])
            _, matched = trie.longest_prefix_match(p, avail, from_root_endpoints=True)
            m = int(matched.cpu()[0]) & 0xFFFFFFFFFFFFFFFF
            assert sorted(e for i, e in enumerate(eps) if m >> i & 1) == op["matched"]
    trie.check_status()


def test_route_in_byte_chunk_mode_vs_oracle():
    """Sequential-equivalent batch routing with 128-BYTE chunks over non-ASCII prompts."""
    import random
    from b200_router.trie import GpuHashTrie
    rng = random.Random(9)
    stems = ["é" * 150, "日本語" * 90, "sys " * 100, "\U0001f600" * 70]
    t = GpuHashTrie(cap_log2=16)
    orc = coracle.TrieOracle(128, by_bytes=True)
    seq = 0
    for b in range(5):
        prompts = [rng.choice(stems) + "".join(rng.choice("abé日") for _ in range(rng.randint(0, 300)))
                   for _ in range(300)]
        avail = 0x3F & ~(1 << 2) if b == 2 else 0x3F
        p = pack_prompts(prompts, chunk_by="bytes")
        ep, depth, matched = t.route(p, avail, 5, seq)
        pk = coracle.Packed([s.encode() for s in prompts])
        oep, odepth, omatched = orc.route_batch(pk.bytes, pk.starts, pk.lens, avail, 5, seq)
        seq += len(prompts)
        assert (ep.cpu().numpy() == oep).all() and (depth.cpu().numpy() == odepth).all()
        assert (u64(matched) == omatched).all()
    t.check_status()
    assert t.dump() == orc.dump()
    t.close()


def test_router_packs_message_parts_without_joining():
    """route_request hands the chat contents to the native packer as parts; the decisions are
    those of routing the "\\n".join(...) strings (routing_logic.py:389-407)."""
    from b200_router import routers

    class Ep:
        def __init__(self, url): self.url = url
    eps = [Ep("http://e%d" % i) for i in range(4)]
    convs = []
    for u in range(40):
        msgs = [{"role": "system", "content": "You are helper %d. " % (u % 5) * 20},
                {"role": "user", "content": "user %d asks é " % (u % 7) * 30}]
        for turn in range(u % 4):
            msgs.append({"role": "assistant", "content": "Hello " * (50 + turn)})
            msgs.append({"role": "user", "content": [{"type": "text", "text": "more %d" % turn},
                                                     {"type": "image_url", "image_url": "x"},
                                                     {"type": "text", "text": "please"}]})
        convs.append({"messages": msgs})
    routers._reset_singletons_for_tests()
    r = routers.PrefixAwareRouter(seed=3)
    got = [asyncio.run(r.route_request(eps, None, None, None, js)) for js in convs]
    routers._reset_singletons_for_tests()
    r2 = routers.PrefixAwareRouter(seed=3)
    want = [r2.route_batch(eps, [routers.extract_prompt(js)])[0] for js in convs]
    routers._reset_singletons_for_tests()
    assert got == want
