"""Canonical tie-break used on both sides of every parity check.

TEST INFRASTRUCTURE (see oracle/__init__.py).

The reference draws `random.choice(list(matched_endpoint))`
(src/vllm_router/routers/routing_logic.py:419).  `list(set_of_str)` order
depends on PYTHONHASHSEED and the draw on the process-global Mersenne Twister,
so the reference is not reproducible across processes.  Parity
is therefore defined with a canonical order (candidates sorted by endpoint
index) and a counter-based draw pick(seed, seq, n), seq = the request's global
sequence number.  `ChoiceShim` plugs that into the unmodified reference by
replacing the `random` name inside `vllm_router.routers.routing_logic`.
"""

M64 = (1 << 64) - 1


def pick(seed: int, seq: int, n: int) -> int:
    """splitmix64 finaliser of seed + seq*golden, mapped to [0, n) by a
    32x32 multiply-high.  Mirrors orc_pick (oracle.c) and b200r_pick (csrc)."""
    x = (seed + seq * 0x9E3779B97F4A7C15) & M64
    x ^= x >> 30
    x = (x * 0xBF58476D1CE4E5B9) & M64
    x ^= x >> 27
    x = (x * 0x94D049BB133111EB) & M64
    x ^= x >> 31
    return ((x >> 32) * n) >> 32


class ChoiceShim:
    """Stand-in for the `random` module seen by the reference router.

    `index_of` maps an endpoint url to its stable index (the bit it occupies in
    the GPU masks); `seq` advances once per routed request, exactly like the
    reference calls random.choice once per request."""

    def __init__(self, index_of, seed: int, seq0: int = 0):
        self.index_of = index_of
        self.seed = seed
        self.seq = seq0

    def choice(self, seq_of_urls):
        cands = sorted(seq_of_urls, key=lambda u: self.index_of[u])
        k = pick(self.seed, self.seq, len(cands))
        self.seq += 1
        return cands[k]
