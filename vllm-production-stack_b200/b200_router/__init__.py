"""b200_router -- B200-native (sm_100a) backend for the vllm_router routing decision.

Public surface (mirrors src/vllm_router/routers/routing_logic.py and
src/vllm_router/prefix/hashtrie.py of the reference):

    RoutingLogic, RoutingInterface, PrefixAwareRouter, SessionRouter,
    initialize_routing_logic, reconfigure_routing_logic, get_routing_logic
    GpuHashTrie (HashTrie), GpuHashRing (uhashring.HashRing subset)
    pack_prompts / PackedPrompts (the packed-bytes-plus-offsets layout)

Everything computes through libb200router.so (hand-written CUDA, C ABI in
include/b200_router.h).  Importing this package does not need a GPU; calling a
router does, and fails loudly without one.
"""
from ._lib import B200RouterError, LIB_PATH, lib  # noqa: F401
from .packing import PackedPrompts, pack_arrays, pack_prompts  # noqa: F401


def __getattr__(name):
    # torch-heavy modules are imported on first use
    if name in ("GpuHashTrie", "decode_dump"):
        from . import trie
        return getattr(trie, name)
    if name == "GpuHashRing":
        from . import ring
        return ring.GpuHashRing
    if name in ("RoutingLogic", "RoutingInterface", "PrefixAwareRouter", "SessionRouter",
                "initialize_routing_logic", "reconfigure_routing_logic", "get_routing_logic",
                "EndpointRegistry", "extract_prompt", "SingletonABCMeta"):
        from . import routers
        return getattr(routers, name)
    raise AttributeError(name)
