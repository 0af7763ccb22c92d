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
    from vllm_router.routers import routing_logic as _ref_rl          # pragma: no cover
    from vllm_router.utils import SingletonABCMeta                    # pragma: no cover
    _RefPrefix, _RefSession = _ref_rl.PrefixAwareRouter, _ref_rl.SessionRouter  # pragma: no cover
    RoutingInterface = _ref_rl.RoutingInterface                       # pragma: no cover
    RoutingLogic = _ref_rl.RoutingLogic                               # pragma: no cover
    HAVE_REFERENCE = True                                             # pragma: no cover
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
    """Stable url -> bit index map feeding the endpoint masks.  Append-only: the
    reference trie keeps stale endpoints forever and filters them by
    intersection with the available set (hashtrie.py:96), so a bit is never
    reused.  EndpointInfo.url is the only identity the hot
    path uses (service_discovery.py:80-110, routing_logic.py:414)."""

    def __init__(self, limit: int = MAX_ENDPOINTS):
        self.limit = limit
        self.urls: List[str] = []
        self.index: Dict[str, int] = {}

    def bit(self, url: str) -> int:
        i = self.index.get(url)
        if i is None:
            if len(self.urls) >= self.limit:
                raise _lib.B200RouterError(
                    "more than %d distinct endpoint urls seen: endpoint masks are one 64-bit "
                    "word" % self.limit)
            i = self.index[url] = len(self.urls)
            self.urls.append(url)
        return i

    def mask(self, endpoints) -> int:
        m = 0
        for e in endpoints:
            m |= 1 << self.bit(e.url if hasattr(e, "url") else e)
        return m


def extract_prompt(request_json: Dict) -> str:
    """Prompt text the reference hashes (routing_logic.py:388-412): chat
    messages' contents joined by newline (text parts of multimodal content
    joined by a blank), else the completions `prompt` (KeyError if absent)."""
    if "messages" in request_json:
        messages = request_json["messages"]
        if not messages:
            return ""
        parts = []
        for message in messages:
            content = message.get("content", "")
            if isinstance(content, list):
                parts.append(" ".join(part.get("text", "") for part in content
                                      if part.get("type") == "text"))
            elif content is not None:
                parts.append(content)
        return "\n".join(parts)
    return request_json["prompt"]


def extract_prompt_parts(request_json: Dict):
    """extract_prompt without the final join: a tuple of parts standing for "\n".join(parts)
    (or the completions prompt itself).  pack_prompts lays the parts end to end with the
    separator between them, so long chat histories are copied once, into the packed buffer,
    instead of twice (routing_logic.py:407 builds the joined string first)."""
    if "messages" in request_json:
        messages = request_json["messages"]
        if not messages:
            return ""
        parts = []
        for message in messages:
            content = message.get("content", "")
            if isinstance(content, list):
                parts.append(" ".join(part.get("text", "") for part in content
                                      if part.get("type") == "text"))
            elif content is not None:
                parts.append(content)
        return tuple(parts)
    return request_json["prompt"]


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
        self.registry = EndpointRegistry()
        self.seed = seed
        self.seq = 0                      # global request sequence number (tie-break counter)
        self.batcher = None
        if micro_batch_ms > 0:
            from .batching import MicroBatcher
            self.batcher = MicroBatcher(self.route_batch, micro_batch_max, micro_batch_ms)
        self._initialized = True

    # ---- batch API
    def route_packed(self, avail_mask: int, packed: PackedPrompts):
        """Route a packed batch; returns device tensors (endpoint idx, depth, matched)."""
        out = self.hashtrie.route(packed, avail_mask, self.seed, self.seq)
        self.seq += packed.n
        return out

    def route_batch(self, endpoints, prompts: Sequence) -> List[str]:
        """Route prompts in list order against one endpoint set; returns urls.  A prompt is a
        str or a tuple of parts (extract_prompt_parts)."""
        if len(prompts) == 0:
            return []
        avail = self.registry.mask(endpoints)
        ep, _, _ = self.route_packed(avail, pack_prompts(prompts, chunk_by=self.chunk_by,
                                                         arena=self._arena))
        idx = ep.cpu().tolist()
        self.hashtrie.check_status()
        if any(i < 0 for i in idx):
            raise IndexError("Cannot choose from an empty sequence")   # random.choice([])
        return [self.registry.urls[i] for i in idx]

    # ---- reference API
    async def route_request(self, endpoints, engine_stats, request_stats, request,
                            request_json) -> str:
        prompt = extract_prompt_parts(request_json)
        if self.batcher is not None:
            return await self.batcher.route(endpoints, prompt)
        return self.route_batch(endpoints, [prompt])[0]


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
        if HAVE_REFERENCE:                                            # pragma: no cover
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
