"""CPU-only checks: the C ABI library loads and exports exactly what the header
declares, and the host-side mirror of the reference interface behaves like the
reference (no compute calls -- those need a GPU and are in test_gpu_*.py)."""
import asyncio
import os
import re
import subprocess

import numpy as np
import pytest

import b200_router
from b200_router import _lib, packing, traces
from b200_router import routers
from oracle import prefix_ref

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "b200_router.h")


def _declared_symbols():
    text = open(HEADER).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(b200r_[a-z0-9_]+)\s*\(", text)))


def test_library_exports_every_declared_symbol():
    lib = _lib.lib()
    declared = _declared_symbols()
    assert len(declared) >= 20
    for name in declared:
        assert hasattr(lib, name), name
    out = subprocess.check_output(["nm", "-D", "--defined-only", _lib.LIB_PATH], text=True)
    exported = sorted(set(re.findall(r"\b(b200r_[a-z0-9_]+)\b", out)))
    assert exported == declared
    # and the ctypes table binds each of them
    assert sorted(_lib.SIGNATURES) == declared


def test_abi_version_and_host_side_pick():
    from oracle.chooser import pick
    lib = _lib.lib()
    assert lib.b200r_abi_version() == 2
    for seed, seq, n in [(0, 0, 1), (7, 123456789, 8), (2**63, 2**40 + 3, 64), (5, 9, 3)]:
        assert lib.b200r_pick(seed, seq, n) == pick(seed, seq, n)


def test_no_cuda_means_loud_failure():
    import torch
    if torch.cuda.is_available():
        pytest.skip("needs a machine without a GPU")
    with pytest.raises(b200_router.B200RouterError):
        routers._reset_singletons_for_tests()
        routers.PrefixAwareRouter()
    with pytest.raises(b200_router.B200RouterError):
        routers.SessionRouter("x-user-id")


def test_invalid_arguments_are_rejected_without_a_device():
    lib = _lib.lib()
    assert lib.b200r_trie_lookup_batch(None, None, None, 1, 0, None, None, None) == -1
    assert b"trie" in lib.b200r_last_error()
    assert lib.b200r_md5_batch(None, None, None, -1, None, None) == -1
    assert lib.b200r_xxh64_chunks_noalloc(None, None, None, None, 1, 128, 0, None, None, 1, None) == -1


def test_pack_prompts_layout():
    prompts = ["abc", "é" * 200, "", "x" * 128, "日本語"]
    p = packing.pack_prompts(prompts, pin=False)
    assert p.n == 5 and not p.all_ascii
    assert (p.starts.numpy() % 16 == 0).all()
    assert p.lens.tolist() == [len(s.encode()) for s in prompts]
    assert p.nchars.tolist() == [len(s) for s in prompts]
    assert p.max_chunks == sum((len(s) + 127) // 128 for s in prompts)
    buf = p.bytes.numpy()
    for s, st, ln in zip(prompts, p.starts.tolist(), p.lens.tolist()):
        assert bytes(buf[st:st + ln]) == s.encode()
    q = packing.pack_prompts(["a" * 300, "b"], pin=False)
    assert q.all_ascii and q.nchars is None and q.max_chunks == 4 and q.payload_bytes == 301


def test_native_packer_exports_its_header_and_equals_the_python_packer():
    """libb200pack.so (csrc/packer.cpp): symbols == include/b200_pack.h, and byte-identical
    output to the pure-Python packer for str prompts, multi-part prompts, both chunking units."""
    import random
    L = packing.pack_lib()
    assert L is not None, "libb200pack.so is not built (python vllm-production-stack_b200/build.py)"
    text = re.sub(r"/\*.*?\*/", "", open(os.path.join(ROOT, "include", "b200_pack.h")).read(), flags=re.S)
    declared = sorted(set(re.findall(r"\b(b200p_[a-z0-9_]+)\s*\(", text)))
    out = subprocess.check_output(["nm", "-D", "--defined-only", packing.PACK_LIB_PATH], text=True)
    assert sorted(set(re.findall(r"\b(b200p_[a-z0-9_]+)\b", out))) == declared and len(declared) == 3
    rng = random.Random(5)
    alphabet = ["a", "b", " ", "\u00e9", "\u65e5", "\U0001f600", "\n"]

    def rand_str(lo, hi):
        return "".join(rng.choice(alphabet) for _ in range(rng.randint(lo, hi)))

    cases = [[], [""], ["", "", ""], ["x" * 16, "y" * 17, "z" * 15], [("a", "b"), (), ("",), ["p", "", "q"]]]
    for _ in range(30):
        batch = []
        for _ in range(rng.randint(1, 40)):
            if rng.random() < 0.4:
                batch.append(tuple(rand_str(0, 300) for _ in range(rng.randint(0, 5))))
            else:
                batch.append(rand_str(0, 700))
        cases.append(batch)
    for batch in cases:
        for by in ("chars", "bytes"):
            for sep in ("\n", " \u00e9 "):
                a = packing._pack_prompts_py(batch, 128, False, by == "bytes", sep)
                b = packing.pack_prompts(batch, pin=False, chunk_by=by, sep=sep)
                assert bytes(a.bytes.numpy()) == bytes(b.bytes.numpy())
                assert a.starts.tolist() == b.starts.tolist() and a.lens.tolist() == b.lens.tolist()
                assert (a.nchars is None) == (b.nchars is None)
                assert a.nchars is None or a.nchars.tolist() == b.nchars.tolist()
                assert (a.n, a.max_chunks, a.all_ascii, a.payload_bytes) == \
                       (b.n, b.max_chunks, b.all_ascii, b.payload_bytes)
    # the multi-threaded copy path (taken for >= 64 pieces): every thread count, ragged sizes
    big = [rand_str(0, 900) if rng.random() < 0.7 else tuple(rand_str(0, 200) for _ in range(rng.randint(0, 6)))
           for _ in range(3000)]
    want = packing._pack_prompts_py(big, 128, False, False, "\n")
    for th in (1, 2, 3, 8, 0):
        got = packing.pack_prompts(big, pin=False, threads=th)
        assert bytes(got.bytes.numpy()) == bytes(want.bytes.numpy()), th
        assert got.starts.tolist() == want.starts.tolist() and got.nchars.tolist() == want.nchars.tolist()
    arena = packing.PackArena(slots=2, pin=False)
    first = packing.pack_prompts(["abc", "de"], pin=False, arena=arena)
    kept = bytes(first.bytes.numpy())
    packing.pack_prompts(["zzzzzzzz"], pin=False, arena=arena)          # other slot
    assert bytes(first.bytes.numpy()) == kept
    pipe = packing.PackPipeline(arena=packing.PackArena(slots=4, pin=False), workers=2, pin=False)
    sets = [["p%d-%d" % (b, i) * (i % 7 + 1) for i in range(50)] for b in range(6)]
    for b in sets[:3]:
        pipe.submit(b)
    with pytest.raises(RuntimeError):       # a 4-slot arena holds 3 queued batches + 1 in use
        pipe.submit(sets[3])
    for k, b in enumerate(sets):            # results come back in submission order
        got = pipe.next()
        assert bytes(got.bytes.numpy()) == bytes(packing._pack_prompts_py(b, 128, False, False, "\n").bytes.numpy())
        if k + 3 < len(sets):
            pipe.submit(sets[k + 3])
    pipe.close()
    with pytest.raises(TypeError):
        packing.pack_prompts(["ok", 3], pin=False)
    with pytest.raises(TypeError):
        packing.pack_prompts([("ok", b"bytes")], pin=False)
    with pytest.raises(ValueError):
        packing.pack_prompts(["x"], chunk_by="words")


def test_byte_chunk_mode_is_the_gateway_pickers_chunking(golden_go):
    """chunk_by="bytes": lens == nchars, i.e. the device chunks 128 BYTES
    (prefix_aware_picker.go:154-165); chunk counts match the golden chunk lists."""
    texts = [bytes.fromhex(c["utf8_hex"]).decode() for c in golden_go["chunks"]]
    p = packing.pack_prompts(texts, pin=False, chunk_by="bytes")
    assert p.all_ascii and p.nchars is None
    assert p.max_chunks == sum(len(c["hashes"]) for c in golden_go["chunks"])
    q = packing.pack_prompts(texts, pin=False)
    assert not q.all_ascii and q.max_chunks < p.max_chunks


def test_extract_prompt_parts_stand_for_the_joined_prompt(golden_prefix):
    tr = [t for t in golden_prefix["traces"] if t["name"] == "request_json_shapes"][0]
    for js in tr["json_inputs"]:
        parts = routers.extract_prompt_parts(js)
        joined = parts if isinstance(parts, str) else "\n".join(parts)
        assert joined == prefix_ref.flatten_prompt(js)
        a = packing.pack_prompts([parts], pin=False)
        b = packing.pack_prompts([joined], pin=False)
        assert bytes(a.bytes.numpy()) == bytes(b.bytes.numpy()) and a.max_chunks == b.max_chunks
    with pytest.raises(KeyError):
        routers.extract_prompt_parts({"neither": 1})


def test_register_tail_hash_equals_xxh64_on_the_host(tmp_path):
    """csrc/tailhash.cuh (the ragged-last-chunk hash of k_prep_scan, written with static
    register indexing) compiled for the host and held against the oracle's XXH64 for every
    length 1..127 on random bytes."""
    from oracle import coracle
    coracle.lib()
    exe = str(tmp_path / "tailhash_host_test")
    odir = os.path.join(ROOT, "oracle")
    subprocess.check_call(["g++", "-O2", "-std=c++17", "-I", os.path.join(ROOT, "vllm-production-stack_b200", "csrc"),
                           os.path.join(ROOT, "tests", "tailhash_host_test.cpp"),
                           os.path.join(odir, "liboracle.so"), "-Wl,-rpath," + odir, "-o", exe])
    out = subprocess.check_output([exe], text=True)
    assert "bad = 0" in out


def test_extract_prompt_matches_reference_shapes(golden_prefix):
    tr = [t for t in golden_prefix["traces"] if t["name"] == "request_json_shapes"][0]
    for js in tr["json_inputs"]:
        assert routers.extract_prompt(js) == prefix_ref.flatten_prompt(js)
    with pytest.raises(KeyError):
        routers.extract_prompt({"neither": 1})


def test_endpoint_registry_keeps_bits_and_recycles_idle_ones():
    reg = routers.EndpointRegistry()

    class Ep:
        def __init__(self, url): self.url = url
    a = reg.mask([Ep("u1"), Ep("u2")])
    b = reg.mask([Ep("u3"), Ep("u1")])
    assert a == 0b011 and b == 0b101 and reg.urls == ["u1", "u2", "u3"]
    # more endpoints at once than mask bits: an error, like before
    small = routers.EndpointRegistry(limit=2)
    small.mask(["a", "b"])
    with pytest.raises(b200_router.B200RouterError):
        small.mask(["a", "b", "c"])
    # pod churn (service_discovery.py:651: a new url per restart): more than `limit` urls over
    # the lifetime.  Bits of urls that are no longer listed are retired (cleared from the
    # table through on_retire) and handed out again; live endpoints keep theirs.
    retired = []
    reg = routers.EndpointRegistry(limit=4, on_retire=retired.append)
    live = ["p0", "p1", "p2"]
    seen_bits = {}
    for gen in range(3, 40):
        m = reg.mask(live)
        assert bin(m).count("1") == 3
        for u in live:
            seen_bits.setdefault(u, reg.index[u])
            assert reg.index[u] == seen_bits[u]          # stable while the endpoint is listed
        live = live[1:] + ["p%d" % gen]                  # oldest pod restarts under a new url
    assert len([u for u in reg.urls if u]) <= 4 and reg.retired >= 30
    assert len(retired) == reg.retired and all(bin(b).count('1') == 1 for b in retired)
    # state for checkpoints: urls in bit order
    want = reg.mask(live)
    st = reg.state()
    reg2 = routers.EndpointRegistry(limit=4)
    reg2.load_state(st)
    assert reg2.mask(live) == want


def test_qps_routing_rule():
    class Ep:
        def __init__(self, url): self.url = url

    class St:
        def __init__(self, qps): self.qps = qps

    class Dummy(routers.RoutingInterface):
        def route_request(self, *a):
            return None
    routers.SingletonABCMeta._instances.pop(Dummy, None)
    d = Dummy()
    eps = [Ep("a"), Ep("b"), Ep("c")]
    assert d._qps_routing(eps, {"a": St(3), "b": St(1), "c": St(1)}) == "b"   # first minimum
    assert d._qps_routing(eps, {"a": St(3), "c": St(1)}) == "b"               # unseen endpoint wins
    routers.SingletonABCMeta._instances.pop(Dummy, None)


def test_registry_names_and_errors():
    assert routers.RoutingLogic("prefixaware") == routers.RoutingLogic.PREFIXAWARE
    assert routers.RoutingLogic("session") == routers.RoutingLogic.SESSION_BASED
    with pytest.raises(ValueError):
        routers.initialize_routing_logic("nonsense")
    routers._reset_singletons_for_tests()
    with pytest.raises(ValueError):
        routers.get_routing_logic()


def test_trace_generators_are_deterministic():
    a = traces.SharedPrefixTrace(n_total=300, batch=100, n_sys=4, sys_chars=256, seed=5)
    b = traces.SharedPrefixTrace(n_total=300, batch=100, n_sys=4, sys_chars=256, seed=5)
    for k in range(3):
        x, y = a.batch_arrays(k), b.batch_arrays(k)
        assert all((u == v).all() for u, v in zip(x, y))
        assert (x[1] % 16 == 0).all()
    p, u = traces.multi_round_qa(users=3, rounds=2, sys_len=5, hist_len=2, answer_len=1)
    assert len(p) == 6 and p[3].startswith(p[0]) and u == [0, 1, 2, 0, 1, 2]
    ids = traces.session_ids_uuid(5, 1)
    assert ids == traces.session_ids_uuid(5, 1) and all(len(i) == 36 and i[14] == "4" for i in ids)


def test_micro_batcher_keeps_arrival_order_and_splits_on_endpoint_change():
    from b200_router.batching import MicroBatcher
    calls = []

    def fake_route_batch(endpoints, prompts):
        calls.append((tuple(endpoints), list(prompts)))
        return ["%s@%s" % (p, endpoints[len(p) % len(endpoints)]) for p in prompts]

    async def go():
        mb = MicroBatcher(fake_route_batch, max_batch=4, max_delay_ms=20)
        eps_a, eps_b = ["u1", "u2"], ["u1", "u3"]
        tasks = [asyncio.ensure_future(mb.route(eps_a, "p%d" % i)) for i in range(6)]      # 4 + 2
        tasks += [asyncio.ensure_future(mb.route(eps_b, "q%d" % i)) for i in range(3)]     # new set
        tasks += [asyncio.ensure_future(mb.route(eps_a, "r0"))]
        return await asyncio.gather(*tasks), mb
    got, mb = asyncio.run(go())
    assert [c[1] for c in calls] == [["p0", "p1", "p2", "p3"], ["p4", "p5"], ["q0", "q1", "q2"], ["r0"]]
    assert [c[0] for c in calls] == [("u1", "u2")] * 2 + [("u1", "u3")] + [("u1", "u2")]
    assert got[0] == "p0@u1" and got[6] == "q0@u1" and got[-1] == "r0@u1"
    assert mb.batches_routed == 4 and mb.requests_routed == 10

    def failing(endpoints, prompts):
        raise KeyError("prompt")

    async def go2():
        mb = MicroBatcher(failing, max_batch=2, max_delay_ms=5)
        return await asyncio.gather(mb.route(["u"], "a"), mb.route(["u"], "b"), return_exceptions=True)
    assert all(isinstance(e, KeyError) for e in asyncio.run(go2()))
