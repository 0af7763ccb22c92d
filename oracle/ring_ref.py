"""Python restatement of the session-routing path.

TEST INFRASTRUCTURE (see oracle/__init__.py).

PARITY UNPINNED for the ring arithmetic: `uhashring` (pinned ==2.3 by the
reference: pyproject.toml:24, requirements.txt:10, uv.lock:1714-1719) is a
third-party dependency that is absent from /root/reference, not installed and
not installable here (no network).  `RingRef` restates its published default
(non-ketama) `MetaRing` behaviour:

  H(k)           = int(md5(str(k).encode("utf-8")).hexdigest(), 16)   (128 bit)
  add_node(n)    : for w in range(160): ring[H(f"{n}-{w}")] = n ; keys = sorted(ring)
  remove_node(n) : delete the same 160 points ; keys = sorted(ring)
  get_node(key)  : None if the ring is empty, else
                   p = bisect.bisect_right(keys, H(key)); p = 0 if p == len(keys); ring[keys[p]]
  get_nodes()    : the node names

The reference's own tests at this boundary (src/tests/test_session_router.py)
assert properties only (stickiness, QPS fallback, "fewer than all remapped"),
never which endpoint a key maps to; tests/test_session_* re-run those
properties against the GPU backend and compare element-wise with this file.
If the real package is importable, tests/test_oracle_golden.py additionally
checks RingRef against it.

SessionRouterRef follows (paths relative to /root/reference):
  RoutingInterface._qps_routing      src/vllm_router/routers/routing_logic.py:60-82
  RoutingInterface._update_hash_ring src/vllm_router/routers/routing_logic.py:84-103
  SessionRouter.route_request        src/vllm_router/routers/routing_logic.py:185-219
"""
import bisect
from hashlib import md5

VNODES = 160


def ring_hash(key) -> int:
    return int(md5(str(key).encode("utf-8")).hexdigest(), 16)


class RingRef:
    def __init__(self, nodes=()):
        self._owner = {}        # 128-bit point -> node name
        self._names = []        # insertion-ordered node names
        self._keys = []         # sorted points
        for n in nodes:
            self.add_node(n)

    def get_nodes(self):
        return list(self._names)

    def add_node(self, name):
        if name in self._names:
            return
        self._names.append(name)
        for w in range(VNODES):
            self._owner[ring_hash("%s-%s" % (name, w))] = name
        self._keys = sorted(self._owner)

    def remove_node(self, name):
        self._names.remove(name)
        for w in range(VNODES):
            del self._owner[ring_hash("%s-%s" % (name, w))]
        self._keys = sorted(self._owner)

    def get_node(self, key):
        if not self._owner:
            return None
        p = bisect.bisect_right(self._keys, ring_hash(key))
        if p == len(self._keys):
            p = 0
        return self._owner[self._keys[p]]

    def points(self):
        """Sorted (point, node name) pairs -- the structure the GPU ring uploads."""
        return [(k, self._owner[k]) for k in self._keys]


def qps_route(endpoints, request_stats):
    """First endpoint (list order) without stats wins, else strictly lowest qps."""
    best = float("inf")
    pick = None
    for ep in endpoints:
        url = ep.url
        if url not in request_stats:
            return url
        q = request_stats[url].qps
        if q < best:
            best = q
            pick = url
    return pick


class SessionRouterRef:
    def __init__(self, session_key):
        if session_key is None:
            raise ValueError("SessionRouter must be initialized with a session_key")
        self.session_key = session_key
        self.ring = RingRef()

    def sync_ring(self, endpoints):
        want = set(ep.url for ep in endpoints)
        have = set(self.ring.get_nodes())
        for gone in have - want:
            self.ring.remove_node(gone)
        for new in want - have:
            self.ring.add_node(new)

    def route_request(self, endpoints, engine_stats, request_stats, request):
        sid = request.headers.get(self.session_key, None)
        self.sync_ring(endpoints)
        if sid is None:
            return qps_route(endpoints, request_stats)
        return self.ring.get_node(sid)
