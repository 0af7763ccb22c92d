#!/usr/bin/env python
"""Generate tests/golden/*.json from the UNMODIFIED reference.

Runs only in the build container (needs /root/reference).  It imports
vllm_router.prefix.hashtrie.HashTrie and
vllm_router.routers.routing_logic.PrefixAwareRouter as they are, with
sys.modules stubs for the two absent third-party packages (`kubernetes`,
pulled in by service_discovery.py:27, and `uhashring`, routing_logic.py:39),
and replaces only the `random` name inside routing_logic by the canonical
chooser (oracle/chooser.py) -- the tie-break is the one thing the reference
cannot reproduce across processes.

Outputs (small, committed):
  tests/golden/xxh64_kat.json     python-xxhash digests (the reference's hash)
  tests/golden/prefix_traces.json decisions, per-request lookup results and the
                                  final trie of the reference for several traces
  tests/golden/md5_ring_kat.json  hashlib md5 digests + the ring restatement's
                                  answers (ring arithmetic itself is unpinned:
                                  uhashring is not available, see oracle/ring_ref.py)
"""
import asyncio
import json
import os
import random
import sys
import types

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF_SRC = "/root/reference/src"
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "vllm-production-stack_b200"))
sys.path.insert(0, REF_SRC)


def _stub_absent_packages():
    for name in ("kubernetes", "kubernetes.client", "kubernetes.config", "kubernetes.watch",
                 "kubernetes.client.rest"):
        if name not in sys.modules:
            sys.modules[name] = types.ModuleType(name)
    k = sys.modules["kubernetes"]
    k.client = sys.modules["kubernetes.client"]
    k.config = sys.modules["kubernetes.config"]
    k.watch = sys.modules["kubernetes.watch"]
    sys.modules["kubernetes.client"].rest = sys.modules["kubernetes.client.rest"]
    sys.modules["kubernetes.client.rest"].ApiException = type("ApiException", (Exception,), {})
    if "uhashring" not in sys.modules:
        m = types.ModuleType("uhashring")
        m.HashRing = type("HashRing", (), {})
        sys.modules["uhashring"] = m


_stub_absent_packages()

import xxhash  # noqa: E402
from vllm_router.prefix.hashtrie import HashTrie  # noqa: E402
import vllm_router.routers.routing_logic as ref_rl  # noqa: E402
from vllm_router.utils import SingletonABCMeta  # noqa: E402

from oracle.chooser import ChoiceShim  # noqa: E402
from oracle.ring_ref import RingRef, ring_hash  # noqa: E402
from b200_router import traces  # noqa: E402


class _Ep:
    def __init__(self, url):
        self.url = url


def dump_ref_trie(trie: HashTrie, index_of):
    out = []
    stack = [((), trie.root)]
    while stack:
        path, node = stack.pop()
        out.append([["%016x" % h for h in path], sorted(index_of[u] for u in node.endpoints)])
        for h, kid in node.children.items():
            stack.append((path + (h,), kid))
    out.sort()
    return out


def run_reference(urls, batches, seed, as_json=None):
    """batches: [(avail idx list, [prompts])].  Returns per-request records."""
    index_of = {u: i for i, u in enumerate(urls)}
    SingletonABCMeta._instances.pop(ref_rl.PrefixAwareRouter, None)
    router = ref_rl.PrefixAwareRouter()
    shim = ChoiceShim(index_of, seed)
    ref_rl.random = shim
    recs = []

    async def go():
        for avail, prompts in batches:
            eps = [_Ep(urls[i]) for i in avail]
            aset = set(urls[i] for i in avail)
            for j, p in enumerate(prompts):
                rj = as_json(p) if as_json else {"prompt": p}
                text = p if not as_json else None
                if text is not None:
                    mlen, mset = await router.hashtrie.longest_prefix_match(text, aset)
                    look = [mlen // 128, sorted(index_of[u] for u in mset)]
                else:
                    look = None
                url = await router.route_request(eps, None, None, None, rj)
                recs.append([index_of[url], look])

    asyncio.run(go())
    state = dump_ref_trie(router.hashtrie, index_of)
    SingletonABCMeta._instances.pop(ref_rl.PrefixAwareRouter, None)
    return recs, state


def gen_xxh64():
    cases = []
    texts = ["", "a", "abc", "1" * 128, "a" * 128, ("hi " * 42 + "hi"), "é" * 128,
             "日本語" * 50, "🙂" * 7 + "x", "\x00" * 40]
    import random
    rng = random.Random(7)
    for n in list(range(0, 70)) + [95, 96, 97, 127, 128, 129, 255, 256, 257, 384, 511, 512, 513]:
        texts.append("".join(chr(rng.randrange(32, 127)) for _ in range(n)))
    for t in texts:
        b = t.encode("utf-8")
        cases.append({"utf8_hex": b.hex(), "xxh64": "%016x" % xxhash.xxh64(t).intdigest()})
    # chunking of whole prompts as the reference's generator does it
    chunked = []
    for t in ["", "x", "1" * 128 + "2" * 128 + "3" * 5, "é" * 300, "aé🙂" * 100, "日本" * 64,
              "".join(chr(rng.randrange(32, 0x250)) for _ in range(700))]:
        hs = ["%016x" % h for h in HashTrie()._chunk_and_hash(t)]
        chunked.append({"text": t, "hashes": hs})
    return {"source": "python-xxhash %s (xxHash %s) + HashTrie._chunk_and_hash, unmodified"
                      % (xxhash.VERSION, xxhash.XXHASH_VERSION),
            "digests": cases, "chunked": chunked}


def gen_prefix():
    out = {"source": "vllm_router PrefixAwareRouter.route_request + HashTrie, unmodified; "
                     "random replaced by oracle.chooser.ChoiceShim", "traces": []}

    def add(name, urls, batches, seed, as_json=None, json_inputs=None):
        recs, state = run_reference(urls, batches, seed, as_json)
        out["traces"].append({
            "name": name, "urls": urls, "seed": seed,
            "batches": [{"avail": a, "prompts": p} for a, p in batches],
            "json_inputs": json_inputs,
            "decisions": [r[0] for r in recs],
            "lookups": [r[1] for r in recs],
            "trie": state,
        })

    # 1. the e2e prompt families (tests/e2e/test-routing.py:299-354), 5x each
    c = 128
    fam = [["1" * c, "1" * c + "2" * c, "1" * c + "2" * c + "3" * c],
           ["2" * c + "3" * c + "4" * c, "2" * c + "3" * c, "2" * c],
           ["5" * c, "5" * c + "6" * c, "5" * c + "6" * c + "7" * c],
           ["8" * c, "8" * c + "9" * c, "8" * c + "9" * c + "10" * c],
           ["11" * c, "11" * c + "12" * c, "11" * c + "12" * c + "13" * c],
           ["14" * c, "14" * c + "15" * c, "14" * c + "15" * c + "16" * c]]
    fail = ["1" * c, "2" * c, "5" * c, "8" * c, "11" * c, "14" * c]
    seq = [s for g in fam for s in g for _ in range(5)] + [s for s in fail for _ in range(5)]
    urls2 = ["http://localhost:8001", "http://localhost:8002"]
    add("e2e_families", urls2, [([0, 1], seq)], 11)

    # 2. heavy sharing + availability churn (several seeds)
    urls6 = traces.endpoint_urls(6)
    for s in range(4):
        add("sharing_stress_%d" % s, urls6,
            traces.sharing_stress(n_batches=6, batch=40, n_endpoints=6, seed=s), 100 + s)

    # 3. mini multi-round QA (C1 shape, scaled down)
    prompts, _ = traces.multi_round_qa(users=12, rounds=4, sys_len=120, hist_len=30,
                                       answer_len=20)
    urls4 = traces.endpoint_urls(4)
    add("multi_round_qa_mini", urls4,
        [([0, 1, 2, 3], prompts[:24]), ([0, 2, 3], prompts[24:36]), ([0, 1, 2, 3], prompts[36:])],
        5)

    # 4. UTF-8 prompts whose chunk boundaries fall inside multi-byte runs, empties
    u = ["", "é" * 128, "é" * 128 + "ü" * 3, "日本語" * 100, "日本語" * 100 + "x", "a" + "🙂" * 200,
         "", "a" * 127 + "é", "a" * 127 + "é" + "b" * 128, "é" * 128, "a" + "🙂" * 200 + "!"]
    add("utf8_mixed", urls4, [([0, 1, 2, 3], u), ([1], u), ([0, 1, 2, 3], u)], 9)

    # 5. request_json shapes (routing_logic.py:388-412)
    jsons = [
        {"messages": [{"role": "user", "content": "A" * 200}]},
        {"messages": [{"role": "user", "content": "A" * 200},
                      {"role": "assistant", "content": "B" * 100},
                      {"role": "user", "content": [{"type": "text", "text": "see"},
                                                   {"type": "image_url", "image_url": {"url": "x"}},
                                                   {"type": "text", "text": "this"}]}]},
        {"messages": []},
        {"messages": [{"role": "user", "content": None}, {"role": "user"}]},
        {"prompt": "A" * 200 + "\n" + "B" * 100},
        {"prompt": ""},
    ]
    recs, state = run_reference(urls4, [([0, 1, 2, 3], list(range(len(jsons))) * 2)], 21,
                                as_json=lambda i: jsons[i])
    out["traces"].append({"name": "request_json_shapes", "urls": urls4, "seed": 21,
                          "batches": [{"avail": [0, 1, 2, 3],
                                       "prompts": list(range(len(jsons))) * 2}],
                          "json_inputs": jsons,
                          "decisions": [r[0] for r in recs], "lookups": None, "trie": state})
    return out


def gen_md5_ring():
    from hashlib import md5
    keys = ["", "a", "abc", "abc123", "http://engine1.com-0", "x" * 55, "y" * 56, "z" * 64,
            "w" * 119, "v" * 120, "é" * 30, "session1"]
    dig = [{"utf8_hex": k.encode().hex(), "md5": md5(k.encode()).hexdigest()} for k in keys]
    ring = RingRef(["http://engine1.com", "http://engine2.com"])
    probe = ["abc123", "session1", "session2", "session3", "session4", "session5"]
    return {"source": "hashlib.md5 (digests); ring answers from oracle/ring_ref.py restatement "
                      "(uhashring unavailable: parity unpinned)",
            "digests": dig,
            "ring_nodes": ["http://engine1.com", "http://engine2.com"],
            "ring_point_0": "%032x" % ring_hash("http://engine1.com-0"),
            "ring_answers": {k: ring.get_node(k) for k in probe}}


def gen_go_picker():
    """Byte-chunk mode: the gateway picker (src/gateway_inference_extension/prefix_aware_picker.go)
    cannot be compiled here (no Go toolchain), so its three trie functions are restated below,
    statement by statement, over python-xxhash (xxhash.Sum64 == XXH64 seed 0):
      chunkAndHash :154-165  slices the Go string by BYTE index, 128 bytes per chunk
      insert       :167-189  endpoint added to the root and to every node of the path
      longestPrefixMatch :191-213  starts from root.endpoints & available (unlike the router,
                             which starts from `available`), stops at a missing child or an
                             empty intersection."""
    import xxhash

    def chunk_and_hash(s: str):
        b = s.encode("utf-8")
        return [xxhash.xxh64(b[i:i + 128]).intdigest() for i in range(0, len(b), 128)]

    class Node:
        def __init__(self):
            self.children, self.endpoints = {}, set()

    root = Node()

    def insert(key, ep):
        node = root
        node.endpoints.add(ep)
        for h in chunk_and_hash(key):
            node = node.children.setdefault(h, Node())
            node.endpoints.add(ep)

    def match(key, avail):
        node = root
        matched = node.endpoints & avail
        for h in chunk_and_hash(key):
            child = node.children.get(h)
            if child is None:
                break
            node = child
            cand = node.endpoints & avail
            if not cand:
                break
            matched = cand
        return matched

    texts = ["", "a", "x" * 128, "y" * 129, "\u00e9" * 64, "\u00e9" * 65, "\u65e5\u672c\u8a9e" * 50,
             "ab\u00e9" * 100, "hi " * 42 + "hi", "\U0001f600" * 33 + "tail"]
    chunks = [{"utf8_hex": t.encode().hex(), "hashes": ["%016x" % h for h in chunk_and_hash(t)]}
              for t in texts]
    rng = random.Random(11)
    eps = ["pod-%d" % i for i in range(5)]
    stems = ["\u00e9" * 70, "sys " * 80, "\u65e5\u672c" * 60]
    ops = []
    for step in range(60):
        key = rng.choice(stems) + "".join(rng.choice("ab\u00e9") for _ in range(rng.randint(0, 200)))
        avail = sorted(rng.sample(eps, rng.randint(1, 5)))
        if step % 3 == 2:
            ops.append({"op": "match", "key": key, "avail": avail, "matched": sorted(match(key, set(avail)))})
        else:
            ep = rng.choice(avail)
            ops.append({"op": "insert", "key": key, "endpoint": ep})
            insert(key, ep)
    return {"source": "restatement of prefix_aware_picker.go:154-213 over python-xxhash %s"
                      % xxhash.VERSION, "endpoints": eps, "chunks": chunks, "ops": ops}


def main():
    gold = os.path.join(ROOT, "tests", "golden")
    os.makedirs(gold, exist_ok=True)
    for name, fn in (("xxh64_kat.json", gen_xxh64), ("prefix_traces.json", gen_prefix),
                     ("md5_ring_kat.json", gen_md5_ring), ("go_picker_kat.json", gen_go_picker)):
        with open(os.path.join(gold, name), "w") as f:
            json.dump(fn(), f, ensure_ascii=True, separators=(",", ":"))
        print(name, os.path.getsize(os.path.join(gold, name)), "bytes")


if __name__ == "__main__":
    main()
