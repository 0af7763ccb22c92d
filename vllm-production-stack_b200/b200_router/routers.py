"""Host-side mirror of the reference routing-logic interface, GPU-backed.

Same names, argument meaning and error behaviour as
src/vllm_router/routers/routing_logic.py (RoutingLogic :50-55, RoutingInterface
:58-124, SessionRouter :170-219, PrefixAwareRouter :347-423, registry functions
:470-526) and src/vllm_router/utils.py:52-65 (SingletonABCMeta), for the two
routing logics on the accelerated path: "prefixaware" and "session".  When the
real `vllm_router` package is importable the GPU routers subclass its
PrefixAwareRouter / SessionRouter, so the caller's
`isinstance(router, (KvawareRouter, PrefixAwareRouter))` dispatch
(services/request_service/request.py:252-261) keeps working; see INTEGRATION.md.

On top of the reference's one-request `route_request` each router has a batch
method (`route_batch`) -- the shape the CUDA kernels are built for.
"""
import abc
import enum
from typing import Dict, List, Optional, Sequence

import numpy as np
import torch

from . import _lib
from .packing import PackArena, PackedPrompts, pack_prompts
from .ring import GpuHashRing
from .trie import GpuHashTrie

try:  # drop-in mode: inherit from the real classes when they are importable
    from vllm_router.routers import routing_logic as _ref_rl
    from vllm_router.utils import SingletonABCMeta          
    _RefPrefix, _RefSession = _ref_rl.PrefixAwareRouter, _ref_rl.SessionRouter
    RoutingInterface = _ref_rl.RoutingInterface             
    RoutingLogic = _ref_rl.RoutingLogic                     
    HAVE_REFERENCE = True                                   
except Exception:
    HAVE_REFERENCE = False

    class SingletonABCMeta(abc.ABCMeta):
        """One instance per class; `create=False` returns None instead of creating."""
        _instances: Dict[type, object] = {}

        def __call__(cls, *args, **kwargs):
            if cls not in cls._instances:
                if kwargs.get("create") is False:
                    return None
                cls._instances[cls] = super().__call__(*args, **kwargs)
            return cls._instances[cls]

    class RoutingLogic(str, enum.Enum):
        ROUND_ROBIN = "roundrobin"
        SESSION_BASED = "session"
        KVAWARE = "kvaware"
        PREFIXAWARE = "prefixaware"
        DISAGGREGATED_PREFILL = "disaggregated_prefill"

    class RoutingInterface(metaclass=SingletonABCMeta):
        def _qps_routing(self, endpoints, request_stats) -> str:
            """First endpoint without stats wins; else strictly lowest qps, first
            minimum in list order (routing_logic.py:60-82)."""
            lowest = float("inf")
            ret = None
            for info in endpoints:
                url = info.url
                if url not in request_stats:
                    return url
                q = request_stats[url].qps
                if q < lowest:
                    lowest = q
                    ret = url
            return ret

        def _update_hash_ring(self, endpoints):
            """Set-difference sync of ring membership (routing_logic.py:84-103)."""
            want = set(e.url for e in endpoints)
            have = set(self.hash_ring.get_nodes())
            for node in have - want:
                self.hash_ring.remove_node(node)
            for node in want - have:
                self.hash_ring.add_node(node)

        @abc.abstractmethod
        def route_request(self, endpoints, engine_stats, request_stats, request) -> str:
            raise NotImplementedError

    _RefPrefix = _RefSession = RoutingInterface


MAX_ENDPOINTS = 64


class EndpointRegistry:
    """Stable url -> bit index map feeding the endpoint masks (EndpointInfo.url is the only
    identity the hot path uses: service_discovery.py:80-110, routing_logic.py:414).

    The reference trie keeps stale endpoints forever and filters them by intersection with
    the available set (hashtrie.py:96), so a bit stays assigned after its endpoint disappears
    (SURVEY.md 7.3 H3).  Masks are one 64-bit word, though, and K8s discovery mints a new url
    per pod restart (service_discovery.py:651): when a 65th url shows up, the bits of urls
    that are not in the current endpoint list are RETIRED -- least recently seen first -- by
    clearing them from every node (`on_retire(bits)`, one sweep kernel) and handed out again.
    Decisions are unaffected unless a retired url comes back (it then starts without history).
    Only more than 64 endpoints available AT ONCE is an error."""

    def __init__(self, limit: int = MAX_ENDPOINTS, on_retire=None):
        self.limit = limit
        self.urls: List[Optional[str]] = []       # bit -> url (None = free)
        self.index: Dict[str, int] = {}
        self.last_seen: Dict[str, int] = {}
        self.on_retire = on_retire
        self.retired = 0
        self._tick = 0
        self._mask_cache = (None, 0)

    def _free_bits(self, keep) -> None:
        """Retire the least recently seen urls that are not in `keep` (half of them, at least
        one, so that the sweep is rare)."""
        idle = sorted((u for u in self.index if u not in keep), key=lambda u: self.last_seen.get(u, 0))
        if not idle:
            raise _lib.B200RouterError(
                "more than %d endpoints are available at once: endpoint masks are one 64-bit "
                "word" % self.limit)
        victims = idle[:max(1, len(idle) // 2)]
        bits = 0
        for u in victims:
            b = self.index.pop(u)
            self.last_seen.pop(u, None)
            self.urls[b] = None
            bits |= 1 << b
        self.retired += len(victims)
        if self.on_retire is not None:
            self.on_retire(bits)

    def bit(self, url: str, keep=()) -> int:
        i = self.index.get(url)
        if i is None:
            if len(self.index) >= self.limit:
                self._free_bits(set(keep) | {url})
            if None in self.urls:
                i = self.urls.index(None)
                self.urls[i] = url
            else:
                i = len(self.urls)
                self.urls.append(url)
            self.index[url] = i
            self._mask_cache = (None, 0)
        return i

    def mask(self, endpoints) -> int:
        urls = tuple(e.url if hasattr(e, "url") else e for e in endpoints)
        if self._mask_cache[0] == urls:
            return self._mask_cache[1]
        self._tick += 1
        m = 0
        for u in urls:
            m |= 1 << self.bit(u, keep=urls)
            self.last_seen[u] = self._tick
        self._mask_cache = (urls, m)
        return m

    def state(self) -> List[str]:
        """urls in bit order ('' = free bit), for checkpoints."""
        return [u or "" for u in self.urls]

    def load_state(self, urls) -> None:
        self.urls = [str(u) or None for u in urls]
        self.index = {u: i for i, u in enumerate(self.urls) if u}
        self.last_seen = {}
        self._mask_cache = (None, 0)


def _message_parts(request_json: Dict):
    """The pieces the reference joins with a newline (routing_logic.py:388-412): one per chat
    message (text parts of multimodal content joined by a blank); None for a completions
    request."""
    if "messages" not in request_json:
        return None
    parts = []
    for message in request_json["messages"] or ():
        content = message.get("content", "")
        if isinstance(content, list):
            parts.append(" ".join(part.get("text", "") for part in content
                                  if part.get("type") == "text"))
        elif content is not None:
            parts.append(content)
    return parts


def extract_prompt(request_json: Dict) -> str:
    """Prompt text the reference hashes (routing_logic.py:388-412): chat messages' contents
    joined by newline, else the completions `prompt` (KeyError if absent)."""
    parts = _message_parts(request_json)
    return request_json["prompt"] if parts is None else "\n".join(parts)


def extract_prompt_parts(request_json: Dict):
    """extract_prompt without the final join: a tuple of parts standing for "\n".join(parts)
    (or the completions prompt itself).  pack_prompts lays the parts end to end with the
    separator between them, so long chat histories are copied once, into the packed buffer,
    instead of twice (routing_logic.py:407 builds the joined string first)."""
    parts = _message_parts(request_json)
    if parts is None:
        return request_json["prompt"]
    return tuple(parts) if parts else ""


class PrefixAwareRouter(_RefPrefix):
    """GPU-backed `prefixaware` routing logic (drop-in for routing_logic.py:347-423)."""

    def __init__(self, cap_log2: int = 22, seed: int = 0, device=None,
                 micro_batch_ms: float = 0.0, micro_batch_max: int = 4096, chunk_by: str = "chars"):
        """micro_batch_ms > 0: concurrent route_request calls arriving within that window
        are routed as one GPU batch, in arrival order (b200_router.batching.MicroBatcher).
        chunk_by: "chars" = 128 code points per chunk (this router, hashtrie.py:55); "bytes" =
        128 UTF-8 bytes per chunk (the gateway picker, prefix_aware_picker.go:154-165)."""
        if hasattr(self, "_initialized"):
            return
        if chunk_by not in ("chars", "bytes"):
            raise ValueError("chunk_by must be 'chars' or 'bytes'")
        self.chunk_by = chunk_by
        self._arena = PackArena(slots=3)
        self.hashtrie = GpuHashTrie(cap_log2=cap_log2, chunk_chars=128, device=device)
        self.registry = EndpointRegistry(on_retire=self.hashtrie.retire_endpoints)
        self.seed = seed
        self.seq = 0                      # global request sequence number (tie-break counter)
        self._res_dev = None              # decisions + latched status word, one D2H per batch
        self._res_host = None
        self._aux = None
        self.batcher = None
        if micro_batch_ms > 0:
            from .batching import MicroBatcher
            self.batcher = MicroBatcher(self.route_batch, micro_batch_max, micro_batch_ms)
        self._initialized = True

    # ---- batch API
    def route_packed(self, avail_mask: int, packed: PackedPrompts, out=None):
        """Route a packed batch; returns device tensors (endpoint idx, depth, matched)."""
        res = self.hashtrie.route(packed, avail_mask, self.seed, self.seq, out=out)
        self.seq += packed.n
        return res

    def _result_buffers(self, n: int):
        if self._res_dev is None or self._res_dev.numel() < n + 1:
            cap = max(256, int((n + 1) * 1.5))
            dev = self.hashtrie.device
            self._res_dev = torch.empty(cap, dtype=torch.int32, device=dev)
            self._res_host = torch.empty(cap, dtype=torch.int32).pin_memory()
            self._aux = (torch.empty(cap, dtype=torch.int32, device=dev),
                         torch.empty(cap, dtype=torch.int64, device=dev))
        return self._res_dev, self._res_host, self._aux

    def route_batch(self, endpoints, prompts: Sequence) -> List[str]:
        """Route prompts in list order against one endpoint set; returns urls.  A prompt is a
        str or a tuple of parts (extract_prompt_parts).  One host synchronisation per batch:
        the decisions and the table's latched status word come back in the same copy."""
        n = len(prompts)
        if n == 0:
            return []
        avail = self.registry.mask(endpoints)
        rd, rh, (depth, matched) = self._result_buffers(n)
        trie = self.hashtrie
        with torch.cuda.device(trie.device):
            self.route_packed(avail, pack_prompts(prompts, chunk_by=self.chunk_by, arena=self._arena),
                              out=(rd[:n], depth, matched))
            trie.status_async(rd[n:n + 1])
            rh[:n + 1].copy_(rd[:n + 1], non_blocking=True)
            torch.cuda.current_stream().synchronize()
        idx = rh[:n].tolist()
        trie.raise_for_status(int(rh[n]) & 0xFFFFFFFF)
        if any(i < 0 for i in idx):
            raise IndexError("Cannot choose from an empty sequence")   # random.choice([])
        urls = self.registry.urls
        return [urls[i] for i in idx]

    # ---- reference API
    async def route_request(self, endpoints, engine_stats, request_stats, request,
                            request_json) -> str:
        prompt = extract_prompt_parts(request_json)
        if self.batcher is not None:
            return await self.batcher.route(endpoints, prompt)
        return self.route_one(endpoints, prompt)

    def route_one(self, endpoints, prompt) -> str:
        """One request, now: the latency path (one kernel launch, b200r_prefix_route_one) when
        the prompt is ASCII, else a batch of one.  Same decision either way."""
        text = prompt if isinstance(prompt, str) else "\n".join(prompt)
        if text.isascii():
            r = self.hashtrie.route_one(text.encode("ascii"), self.registry.mask(endpoints),
                                        self.seed, self.seq)
            if r is not None:
                self.seq += 1
                self.hashtrie.raise_for_status(r[3])
                if r[0] < 0:
                    raise IndexError("Cannot choose from an empty sequence")   # random.choice([])
                return self.registry.urls[r[0]]
        return self.route_batch(endpoints, [prompt])[0]

    # ---- checkpoint (the reference has none: its trie dies with the process)
    def save(self, path):
        """Table + what gives its masks a meaning: the url of every mask bit, the chunking
        mode, the tie-break counter."""
        self.hashtrie.save(path, extra={"urls": np.array(self.registry.state(), dtype=np.str_),
                                        "chunk_by": np.array(self.chunk_by),
                                        "seq": np.int64(self.seq), "seed": np.int64(self.seed)})

    def load(self, path, cap_log2: Optional[int] = None):
        """Replace this router's table by a saved one (same process-wide singleton)."""
        trie, extra = GpuHashTrie.load(path, cap_log2=cap_log2 or self.hashtrie.cap_log2,
                                       device=self.hashtrie.device, with_extra=True)
        if str(extra["chunk_by"]) != self.chunk_by:
            trie.close()
            raise _lib.B200RouterError("checkpoint was written with chunk_by=%s" % extra["chunk_by"])
        self.hashtrie.close()
        self.hashtrie = trie
        self.registry = EndpointRegistry(on_retire=trie.retire_endpoints)
        self.registry.load_state([str(u) for u in extra["urls"]])
        self.seq = int(extra["seq"])
        self.seed = int(extra["seed"])


class SessionRouter(_RefSession):
    """GPU-backed `session` routing logic (drop-in for routing_logic.py:170-219)."""

    def __init__(self, session_key: str = None, device=None):
        if hasattr(self, "_initialized"):
            return
        if session_key is None:
            raise ValueError("SessionRouter must be initialized with a session_key")
        self.session_key = session_key
        self.hash_ring = GpuHashRing(device=device)
        self._initialized = True

    def route_batch(self, endpoints, request_stats, session_ids: Sequence[Optional[str]]) -> List[str]:
        """Batch of route_request for one endpoint list; None ids take the QPS fallback."""
        self._update_hash_ring(endpoints)
        keyed = [(i, s) for i, s in enumerate(session_ids) if s is not None]
        out: List[Optional[str]] = [None] * len(session_ids)
        if keyed:
            urls = self.hash_ring.get_node_batch([s for _, s in keyed])
            for (i, _), u in zip(keyed, urls):
                out[i] = u
        for i, s in enumerate(session_ids):
            if s is None:
                out[i] = self._qps_routing(endpoints, request_stats)
        return out

    def route_request(self, endpoints, engine_stats, request_stats, request) -> str:
        session_id = request.headers.get(self.session_key, None)
        self._update_hash_ring(endpoints)
        if session_id is None:
            return self._qps_routing(endpoints, request_stats)
        return self.hash_ring.get_node(session_id)


_OUT_OF_SCOPE = ("roundrobin", "kvaware", "disaggregated_prefill")


def initialize_routing_logic(routing_logic, *args, **kwargs):
    """routing_logic.py:470-497 for the two GPU-backed logics."""
    rl = RoutingLogic(routing_logic) if not isinstance(routing_logic, RoutingLogic) else routing_logic
    if rl == RoutingLogic.SESSION_BASED:
        return SessionRouter(kwargs.get("session_key"))
    if rl == RoutingLogic.PREFIXAWARE:
        return PrefixAwareRouter()
    if rl.value in _OUT_OF_SCOPE:
        if HAVE_REFERENCE:                                  
            return _ref_rl.initialize_routing_logic(rl, *args, **kwargs)
        raise ValueError("routing logic %r is not on the GPU path; use the reference router" % rl.value)
    raise ValueError(f"Invalid routing logic {routing_logic}")


def reconfigure_routing_logic(routing_logic, *args, **kwargs):
    """routing_logic.py:500-512: drops Session (ring rebuilt lazily) but keeps
    the prefix trie across a dynamic reconfigure."""
    inst = SingletonABCMeta._instances.pop(SessionRouter, None)
    if inst is not None:
        inst.hash_ring.close()
    return initialize_routing_logic(routing_logic, *args, **kwargs)


def get_routing_logic():
    """routing_logic.py:515-526 (lookup order: Session before Prefix)."""
    for cls in (SessionRouter, PrefixAwareRouter):
        if cls in SingletonABCMeta._instances:
            return cls()
    raise ValueError("The global router has not been initialized")


def _reset_singletons_for_tests():
    for cls in (SessionRouter, PrefixAwareRouter):
        inst = SingletonABCMeta._instances.pop(cls, None)
        if inst is not None:
            if hasattr(inst, "hashtrie"):
                inst.hashtrie.close()
            if hasattr(inst, "hash_ring"):
                inst.hash_ring.close()
