"""Parity of the CUDA session-routing path (md5 + ring) with the oracle."""
import hashlib

import numpy as np
import pytest

from oracle import coracle, ring_ref
from b200_router import traces
from b200_router.packing import pack_arrays, pack_prompts

from helpers import tight_pack

pytestmark = pytest.mark.gpu


@pytest.fixture()
def ring():
    from b200_router.ring import GpuHashRing
    r = GpuHashRing()
    yield r
    r.close()


def test_md5_golden_and_lengths(ring, golden_md5):
    keys = [bytes.fromhex(c["utf8_hex"]).decode("utf-8") for c in golden_md5["digests"]]
    got = ring.md5_ints(keys)
    assert ["%032x" % g for g in got] == [c["md5"] for c in golden_md5["digests"]]
    rng = np.random.default_rng(2)
    keys = ["".join(map(chr, rng.integers(32, 127, size=n))) for n in range(0, 260)]
    assert ring.md5_ints(keys) == [int(hashlib.md5(k.encode()).hexdigest(), 16) for k in keys]
    # unaligned starts
    out = ring.md5_packed(tight_pack(keys).to(ring.device)).cpu().numpy().view(np.uint64)
    assert [(int(out[2 * i]) << 64) | int(out[2 * i + 1]) for i in range(len(keys))] == \
        [int(hashlib.md5(k.encode()).hexdigest(), 16) for k in keys]


def test_ring_matches_restatement(ring, golden_md5):
    ref = ring_ref.RingRef()
    for n in golden_md5["ring_nodes"]:
        ring.add_node(n)
        ref.add_node(n)
    for k, v in golden_md5["ring_answers"].items():
        assert ring.get_node(k) == v
    nodes = traces.endpoint_urls(16)
    for n in nodes:
        ring.add_node(n)
        ref.add_node(n)
    keys = traces.session_ids_uuid(5000, 1) + [str(i) for i in range(1, 3000)] + ["", "é" * 40, "k" * 200]
    assert ring.get_node_batch(keys) == [ref.get_node(k) for k in keys]
    for gone in (nodes[3], golden_md5["ring_nodes"][0], nodes[15]):
        ring.remove_node(gone)
        ref.remove_node(gone)
        assert ring.get_node_batch(keys) == [ref.get_node(k) for k in keys]
    ring.add_node(nodes[3])
    ref.add_node(nodes[3])
    assert ring.get_node_batch(keys) == [ref.get_node(k) for k in keys]
    assert sorted(ring.get_nodes()) == sorted(ref.get_nodes())


def test_ring_1m_ids_vs_c_oracle(ring):
    """BASELINE config C2 at full size: 1M uuid-shaped ids over 16 endpoints."""
    nodes = traces.endpoint_urls(16)
    ref = ring_ref.RingRef(nodes)
    for n in nodes:
        ring.add_node(n)
    pts = ref.points()
    hi = np.array([p >> 64 for p, _ in pts], dtype=np.uint64)
    lo = np.array([p & (2**64 - 1) for p, _ in pts], dtype=np.uint64)
    owner = np.array([nodes.index(nm) for _, nm in pts], dtype=np.int32)
    buf, starts, lens = traces.session_ids_packed(1 << 20, seed=0)
    want = coracle.ring_lookup_batch(hi, lo, owner, buf, starts, lens)
    got = ring.lookup_packed(pack_arrays(buf, starts, lens, pin=False)).cpu().numpy()
    names = ring.owner_names()
    assert [names[i] for i in got[:1000]] == [nodes[i] for i in want[:1000]]
    remap = np.array([nodes.index(nm) for nm in names], dtype=np.int32)
    assert (remap[got] == want).all()
    # every endpoint gets a fair share (ring sanity, not a parity property)
    share = np.bincount(want, minlength=16) / len(want)
    assert share.min() > 0.03 and share.max() < 0.10


def test_empty_ring_returns_none(ring):
    assert ring.get_node("abc") is None
    got = ring.lookup_packed(pack_prompts(["a", "b"], pin=False)).cpu().tolist()
    assert got == [-1, -1]


class Ep:
    def __init__(self, url): self.url = url


class St:
    def __init__(self, qps): self.qps = qps


class Rq:
    def __init__(self, headers): self.headers = headers


def _fresh_router():
    from b200_router import routers
    routers._reset_singletons_for_tests()
    return routers, routers.initialize_routing_logic("session", session_key="session_id")


def test_session_router_reference_properties():
    """The reference's own unit tests (src/tests/test_session_router.py:24-260)
    against the GPU backend, plus element-wise parity with the restatement."""
    routers, r = _fresh_router()
    assert routers.get_routing_logic() is r and r.session_key == "session_id"
    eps = [Ep("http://engine1.com"), Ep("http://engine2.com")]
    stats = {"http://engine1.com": St(10), "http://engine2.com": St(5)}
    req = Rq({"session_id": "abc123"})
    url = r.route_request(eps, None, stats, req)
    assert url == r.route_request(eps, None, stats, req)
    assert r.route_request(eps, None, stats, Rq({})) == "http://engine2.com"
    # dynamic endpoints
    eps.append(Ep("http://engine3.com"))
    stats["http://engine3.com"] = St(2)
    assert r.route_request(eps, None, stats, req) in [e.url for e in eps]
    # remove a node: valid urls, fewer than all remapped; parity with the restatement
    ref = ring_ref.SessionRouterRef("session_id")
    sids = ["session1", "session2", "session3", "session4", "session5"]
    reqs = [Rq({"session_id": s}) for s in sids]
    before = [r.route_request(eps, None, stats, q) for q in reqs]
    assert before == [ref.route_request(eps, None, stats, q) for q in reqs]
    gone = eps.pop(1)
    del stats[gone.url]
    after = [r.route_request(eps, None, stats, q) for q in reqs]
    assert after == [ref.route_request(eps, None, stats, q) for q in reqs]
    assert all(u in [e.url for e in eps] for u in after)
    assert sum(a != b for a, b in zip(before, after)) < len(sids)
    # add one
    eps.append(Ep("http://engine4.com"))
    stats["http://engine4.com"] = St(1)
    after2 = [r.route_request(eps, None, stats, q) for q in reqs]
    assert after2 == [ref.route_request(eps, None, stats, q) for q in reqs]
    assert sum(a != b for a, b in zip(after, after2)) < len(sids)
    # batch API incl. header-less requests
    got = r.route_batch(eps, stats, ["session1", None, "abc123"])
    assert got == [ref.route_request(eps, None, stats, Rq({"session_id": "session1"})),
                   ref.route_request(eps, None, stats, Rq({})),
                   ref.route_request(eps, None, stats, Rq({"session_id": "abc123"}))]
    with pytest.raises(ValueError):
        routers._reset_singletons_for_tests()
        routers.SessionRouter(None)
    routers._reset_singletons_for_tests()
