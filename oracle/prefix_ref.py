"""Python restatement of the reference prefix-aware routing path.

TEST INFRASTRUCTURE (see oracle/__init__.py).  Used (a) as a second,
independent checker next to oracle.c and (b) as the timed "reference arm" /
cpu_baseline of bench.py: the reference is pure Python, cannot travel to the
GPU box, and its cost is dominated by interpreter work (asyncio lock
round-trips, generator, set copies -- SURVEY.md 3.2), so the restatement keeps
the same asyncio structure: one never-contended lock per node, taken twice per
chunk in both passes, and the prompt hashed twice (lookup, then insert).

Follows (paths relative to /root/reference):
  TrieNode                        src/vllm_router/prefix/hashtrie.py:24-32
  HashTrie._chunk_and_hash        src/vllm_router/prefix/hashtrie.py:45-56
  HashTrie.insert                 src/vllm_router/prefix/hashtrie.py:58-74
  HashTrie.longest_prefix_match   src/vllm_router/prefix/hashtrie.py:76-103
  PrefixAwareRouter.route_request src/vllm_router/routers/routing_logic.py:363-423
Pinned by tests/golden/prefix_*.json (generated from the unmodified reference
by scripts/gen_golden.py).
"""
import asyncio

try:  # the reference calls python-xxhash (C); same wheel is in this image
    import xxhash

    def _xxh64(text: str) -> int:
        return xxhash.xxh64(text).intdigest()
except ImportError:  # pragma: no cover - fall back to the C restatement
    from . import coracle

    def _xxh64(text: str) -> int:
        return coracle.xxh64(text.encode("utf-8"))


class _Node:
    __slots__ = ("kids", "eps", "gate")

    def __init__(self):
        self.kids = {}          # chunk hash -> _Node
        self.eps = set()        # endpoint urls routed through this node
        self.gate = asyncio.Lock()


class PrefixTrieRef:
    """Trie keyed by per-chunk xxh64 with one endpoint set per node."""

    def __init__(self, chunk_chars: int = 128):
        self.top = _Node()
        self.chunk_chars = chunk_chars

    def chunk_hashes(self, text: str):
        step = self.chunk_chars
        for off in range(0, len(text), step):
            yield _xxh64(text[off:off + step])

    async def add(self, text: str, endpoint: str) -> None:
        cur = self.top
        async with cur.gate:
            cur.eps.add(endpoint)
        for h in self.chunk_hashes(text):
            async with cur.gate:
                nxt = cur.kids.get(h)
                if nxt is None:
                    nxt = cur.kids[h] = _Node()
                cur = nxt
            async with cur.gate:
                cur.eps.add(endpoint)

    async def match(self, text: str, available: set):
        cur = self.top
        matched_chars = 0
        chosen = available
        for h in self.chunk_hashes(text):
            async with cur.gate:
                cur = cur.kids.get(h)
            if not cur:
                break
            async with cur.gate:
                here = cur.eps.copy()
            both = here.intersection(chosen)
            if not both:
                break
            matched_chars += self.chunk_chars
            chosen = both
        return matched_chars, chosen

    # -- state export for parity checks -------------------------------------
    def dump(self):
        """{tuple(chunk hashes from the root): frozenset(urls)}; () is the root."""
        out = {}
        stack = [((), self.top)]
        while stack:
            path, node = stack.pop()
            out[path] = frozenset(node.eps)
            for h, kid in node.kids.items():
                stack.append((path + (h,), kid))
        return out


def flatten_prompt(request_json: dict) -> str:
    """routing_logic.py:388-412: chat messages joined by newline (multimodal
    text parts joined by a blank), else the completions `prompt` field."""
    if "messages" in request_json:
        msgs = request_json["messages"]
        if not msgs:
            return ""
        parts = []
        for m in msgs:
            body = m.get("content", "")
            if isinstance(body, list):
                parts.append(" ".join(p.get("text", "") for p in body if p.get("type") == "text"))
            elif body is not None:
                parts.append(body)
        return "\n".join(parts)
    return request_json["prompt"]


class PrefixRouterRef:
    """Sequential router: lookup, tie-break, insert -- one request at a time."""

    def __init__(self, chooser, chunk_chars: int = 128):
        self.trie = PrefixTrieRef(chunk_chars)
        self.chooser = chooser          # object with .choice(seq), e.g. ChoiceShim

    async def route(self, endpoint_urls, request_json) -> str:
        prompt = flatten_prompt(request_json)
        available = set(endpoint_urls)
        _, matched = await self.trie.match(prompt, available)
        url = self.chooser.choice(list(matched))
        await self.trie.add(prompt, url)
        return url

    async def route_prompt(self, available: set, prompt: str) -> str:
        _, matched = await self.trie.match(prompt, available)
        url = self.chooser.choice(list(matched))
        await self.trie.add(prompt, url)
        return url

    def route_many(self, endpoint_urls, prompts):
        """Drive one event loop over a list of prompts (arrival order)."""
        available = set(endpoint_urls)

        async def run():
            return [await self.route_prompt(available, p) for p in prompts]

        return asyncio.run(run())
