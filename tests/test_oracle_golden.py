"""The oracle against the reference's behaviour (CPU only).

Fixtures in tests/golden/ were produced by scripts/gen_golden.py from the
unmodified reference (HashTrie, PrefixAwareRouter.route_request, python-xxhash).
"""
import hashlib

import numpy as np
import pytest

from oracle import coracle, prefix_ref, ring_ref
from oracle.chooser import ChoiceShim, pick
from b200_router import traces


def test_xxh64_known_answers():
    # SURVEY.md 8c known answers (seed 0)
    kat = {b"": 0xef46db3751d8e999, b"a": 0xd24ec4f1a98c6e5b, b"abc": 0x44bc2cf5ad770999,
           b"1" * 128: 0x2ceab0ab9ef63228, b"a" * 128: 0x6550288941193272,
           ("hi " * 42 + "hi").encode(): 0x380de794396a43ea,
           ("é" * 128).encode(): 0x9b2210ff98cf44b0}
    for k, v in kat.items():
        assert coracle.xxh64(k) == v


def test_xxh64_golden_digests(golden_xxh):
    for c in golden_xxh["digests"]:
        assert "%016x" % coracle.xxh64(bytes.fromhex(c["utf8_hex"])) == c["xxh64"]


def test_xxh64_library_if_present():
    xxhash = pytest.importorskip("xxhash")
    rng = np.random.default_rng(0)
    for n in list(range(0, 200)) + [1000, 4097]:
        b = rng.integers(0, 256, size=n, dtype=np.uint8).tobytes()
        assert coracle.xxh64(b) == xxhash.xxh64(b).intdigest()


def test_chunking_golden(golden_xxh):
    for c in golden_xxh["chunked"]:
        hs = coracle.xxh64_chunks(c["text"].encode("utf-8"), 128)
        assert ["%016x" % int(h) for h in hs] == c["hashes"]
        t = prefix_ref.PrefixTrieRef()
        assert ["%016x" % h for h in t.chunk_hashes(c["text"])] == c["hashes"]


def test_pick_same_in_c_and_python():
    for seed in (0, 1, 2**63 + 5):
        for seq in (0, 1, 77, 2**40):
            for n in (1, 2, 3, 8, 64):
                assert coracle.pick(seed, seq, n) == pick(seed, seq, n) < n


def _trace_prompts(tr, batch):
    if tr["json_inputs"] is not None:
        return [prefix_ref.flatten_prompt(tr["json_inputs"][i]) for i in batch["prompts"]]
    return batch["prompts"]


def test_c_trie_matches_reference(golden_prefix):
    for tr in golden_prefix["traces"]:
        t = coracle.TrieOracle()
        seq = 0
        dec, looks = [], []
        for b in tr["batches"]:
            prompts = _trace_prompts(tr, b)
            avail = sum(1 << i for i in b["avail"])
            buf, starts, lens, _ = traces.pack_strings(prompts)
            ep, depth, matched = t.route_batch(buf, starts, lens, avail, tr["seed"], seq)
            seq += len(prompts)
            dec += [int(e) for e in ep]
            looks += [[int(d), [i for i in range(64) if (int(m) >> i) & 1]]
                      for d, m in zip(depth, matched)]
        assert dec == tr["decisions"], tr["name"]
        if tr["lookups"] is not None:
            assert looks == tr["lookups"], tr["name"]
        got = sorted([["%016x" % h for h in path], [i for i in range(64) if (m >> i) & 1]]
                     for path, m in t.dump().items())
        assert got == tr["trie"], tr["name"]


def test_python_restatement_matches_reference(golden_prefix):
    for tr in golden_prefix["traces"]:
        urls = tr["urls"]
        index_of = {u: i for i, u in enumerate(urls)}
        r = prefix_ref.PrefixRouterRef(ChoiceShim(index_of, tr["seed"]))
        dec = []
        for b in tr["batches"]:
            prompts = _trace_prompts(tr, b)
            got = r.route_many([urls[i] for i in b["avail"]], prompts)
            dec += [index_of[u] for u in got]
        assert dec == tr["decisions"], tr["name"]
        got = sorted([["%016x" % h for h in path], sorted(index_of[u] for u in eps)]
                     for path, eps in r.trie.dump().items())
        assert got == tr["trie"], tr["name"]


def test_md5_golden(golden_md5):
    for c in golden_md5["digests"]:
        b = bytes.fromhex(c["utf8_hex"])
        assert coracle.md5(b).hex() == c["md5"] == hashlib.md5(b).hexdigest()
    rng = np.random.default_rng(1)
    for n in range(0, 200):
        b = rng.integers(0, 256, size=n, dtype=np.uint8).tobytes()
        assert coracle.md5(b) == hashlib.md5(b).digest()


def test_ring_restatement(golden_md5):
    ring = ring_ref.RingRef(golden_md5["ring_nodes"])
    assert "%032x" % ring_ref.ring_hash("http://engine1.com-0") == golden_md5["ring_point_0"]
    # SURVEY.md 8c values
    assert golden_md5["ring_point_0"] == "12315b0a9f6f263d54b8f3d5a1f770e4"
    assert "%032x" % ring_ref.ring_hash("abc123") == "e99a18c428cb38d5f260853678922e03"
    for k, v in golden_md5["ring_answers"].items():
        assert ring.get_node(k) == v
    assert len(ring.points()) == 320


def test_ring_against_real_uhashring_if_present():
    uhashring = pytest.importorskip("uhashring")
    nodes = traces.endpoint_urls(16)
    real = uhashring.HashRing()
    mine = ring_ref.RingRef()
    for n in nodes:
        real.add_node(n)
        mine.add_node(n)
    keys = traces.session_ids_uuid(2000, 3)
    assert [real.get_node(k) for k in keys] == [mine.get_node(k) for k in keys]
    real.remove_node(nodes[3])
    mine.remove_node(nodes[3])
    assert [real.get_node(k) for k in keys] == [mine.get_node(k) for k in keys]


def test_c_ring_lookup_matches_python():
    nodes = traces.endpoint_urls(16)
    ring = ring_ref.RingRef(nodes)
    pts = ring.points()
    hi = np.array([p >> 64 for p, _ in pts], dtype=np.uint64)
    lo = np.array([p & (2**64 - 1) for p, _ in pts], dtype=np.uint64)
    owner = np.array([nodes.index(n) for _, n in pts], dtype=np.int32)
    keys = traces.session_ids_uuid(3000, 5) + ["", "1", "12345", "k" * 70]
    buf, starts, lens, _ = traces.pack_strings(keys)
    got = coracle.ring_lookup_batch(hi, lo, owner, buf, starts, lens)
    assert [nodes[i] for i in got] == [ring.get_node(k) for k in keys]


def test_session_ref_properties():
    # the reference's own property tests (src/tests/test_session_router.py:24-63)
    class Ep:
        def __init__(self, url): self.url = url

    class St:
        def __init__(self, qps): self.qps = qps

    class Rq:
        def __init__(self, headers): self.headers = headers

    eps = [Ep("http://engine1.com"), Ep("http://engine2.com")]
    stats = {"http://engine1.com": St(10), "http://engine2.com": St(5)}
    r = ring_ref.SessionRouterRef("session_id")
    a = r.route_request(eps, None, stats, Rq({"session_id": "abc123"}))
    assert a == r.route_request(eps, None, stats, Rq({"session_id": "abc123"}))
    assert r.route_request(eps, None, stats, Rq({})) == "http://engine2.com"


def test_go_picker_byte_chunks_and_trie(golden_go):
    """Byte-chunk mode of the C oracle against the restated gateway picker
    (scripts/gen_golden.py:gen_go_picker; prefix_aware_picker.go:154-213): chunk hashes, then
    the insert / longestPrefixMatch trace (match starts from root.endpoints & available)."""
    for c in golden_go["chunks"]:
        data = bytes.fromhex(c["utf8_hex"])
        got = coracle.xxh64_chunks(data, 128, by_bytes=True)
        assert ["%016x" % int(h) for h in got] == c["hashes"]
    eps = golden_go["endpoints"]
    t = coracle.TrieOracle(128, by_bytes=True)
    for op in golden_go["ops"]:
        pk = coracle.Packed([op["key"].encode()])
        if op["op"] == "insert":
            t.insert_batch(pk.bytes, pk.starts, pk.lens, np.array([eps.index(op["endpoint"])], dtype=np.int32))
        else:
            avail = sum(1 << eps.index(e) for e in op["avail"])
            depth, matched = t.lookup_batch(pk.bytes, pk.starts, pk.lens, avail)
            m = int(matched[0]) if depth[0] > 0 else (t.dump()[()] & avail)
            assert sorted(e for i, e in enumerate(eps) if m >> i & 1) == op["matched"]
