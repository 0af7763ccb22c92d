"""Drop-in of the GPU routers into the UNMODIFIED reference package (CPU test, no GPU).

The reference's routing_logic.py is loaded from /root/reference/src with the few lines of
INTEGRATION.md section 1 appended (in memory; the reference tree is never written), with the
two absent third-party packages stubbed exactly like scripts/gen_golden.py does, and with the
GPU objects (table, ring) replaced by inert stand-ins -- what is under test is the class /
registry / singleton wiring the caller relies on:
  routing_logic.py:470-526   initialize_ / reconfigure_ / get_routing_logic
  request.py:252-272         isinstance dispatch (async 5-argument vs sync 4-argument), session_key
  utils.py:52-65             SingletonABCMeta
Runs in a subprocess so that the patched modules do not leak into the other tests.  Skipped
where the reference tree is not present (the GPU box)."""
import os
import subprocess
import sys
import textwrap

import pytest

REF_SRC = "/root/reference/src"
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

SCRIPT = r'''
import asyncio, inspect, os, re, sys, types
ROOT, REF_SRC = sys.argv[1], sys.argv[2]
sys.path.insert(0, os.path.join(ROOT, "vllm-production-stack_b200"))
sys.path.insert(0, REF_SRC)
for name in ("kubernetes", "kubernetes.client", "kubernetes.config", "kubernetes.watch", "kubernetes.client.rest"):
    sys.modules.setdefault(name, types.ModuleType(name))
k = sys.modules["kubernetes"]
k.client, k.config, k.watch = (sys.modules["kubernetes." + n] for n in ("client", "config", "watch"))
sys.modules["kubernetes.client"].rest = sys.modules["kubernetes.client.rest"]
sys.modules["kubernetes.client.rest"].ApiException = type("ApiException", (Exception,), {})
uh = types.ModuleType("uhashring"); uh.HashRing = type("HashRing", (), {}); sys.modules["uhashring"] = uh

# inert stand-ins for the device objects (constructed by the GPU routers)
import b200_router.trie as bt, b200_router.ring as br
class FakeTrie:
    made = 0
    def __init__(self, *a, **kw): FakeTrie.made += 1; self.device = "cpu"; self.cap_log2 = 22
    def retire_endpoints(self, bits): pass
    def close(self): pass
class FakeRing:
    def __init__(self, *a, **kw): self.nodes = []
    def get_nodes(self): return list(self.nodes)
    def add_node(self, n): self.nodes.append(n)
    def remove_node(self, n): self.nodes.remove(n)
    def get_node(self, key): return self.nodes[len(str(key)) % len(self.nodes)] if self.nodes else None
    def close(self): pass
bt.GpuHashTrie, br.GpuHashRing = FakeTrie, FakeRing
import b200_router.packing as bp
bp.PackArena = lambda *a, **kw: None

# the reference module + the INTEGRATION.md patch, executed as vllm_router.routers.routing_logic
import vllm_router.routers                                    # the package itself, unmodified
path = os.path.join(REF_SRC, "vllm_router", "routers", "routing_logic.py")
src = open(path).read()
integration = open(os.path.join(ROOT, "INTEGRATION.md")).read()
patch = re.search(r"```python\n(try:.*?from b200_router\.routers import.*?)```", integration, re.S).group(1)
marker = "\n# Instead of managing a global _global_router"
assert marker in src, "reference layout changed: cannot place the patch after the class definitions"
src = src.replace(marker, "\n" + patch + marker, 1)
mod = types.ModuleType("vllm_router.routers.routing_logic")
mod.__file__ = path
mod.__package__ = "vllm_router.routers"
sys.modules["vllm_router.routers.routing_logic"] = mod
vllm_router.routers.routing_logic = mod
exec(compile(src, path, "exec"), mod.__dict__)
rl = mod
import b200_router.routers as gpu
assert gpu.HAVE_REFERENCE
# the names in the reference module now are the GPU classes, and those subclass the originals
assert rl.PrefixAwareRouter is gpu.PrefixAwareRouter and rl.SessionRouter is gpu.SessionRouter
orig_prefix = [b for b in gpu.PrefixAwareRouter.__mro__ if b.__name__ == "PrefixAwareRouter" and b is not gpu.PrefixAwareRouter]
orig_session = [b for b in gpu.SessionRouter.__mro__ if b.__name__ == "SessionRouter" and b is not gpu.SessionRouter]
assert orig_prefix and orig_prefix[0].__module__ == "vllm_router.routers.routing_logic"
assert orig_session and issubclass(gpu.SessionRouter, rl.RoutingInterface)
assert type(gpu.PrefixAwareRouter) is rl.SingletonABCMeta

# registry (routing_logic.py:470-526), through the REFERENCE's own functions
r = rl.initialize_routing_logic(rl.RoutingLogic.PREFIXAWARE)
assert type(r) is gpu.PrefixAwareRouter and rl.get_routing_logic() is r
assert gpu.PrefixAwareRouter() is r and FakeTrie.made == 1          # singleton, __init__ guarded
s = rl.initialize_routing_logic(rl.RoutingLogic.SESSION_BASED, session_key="x-user-id")
assert type(s) is gpu.SessionRouter and s.session_key == "x-user-id"
assert rl.get_routing_logic() is s                                    # Session is looked up first
# request.py:252-261: the caller's dispatch
assert isinstance(r, (rl.KvawareRouter, rl.PrefixAwareRouter))
assert not isinstance(s, (rl.KvawareRouter, rl.PrefixAwareRouter))
assert inspect.iscoroutinefunction(r.route_request) and not inspect.iscoroutinefunction(s.route_request)
assert len(inspect.signature(r.route_request).parameters) == 5
assert len(inspect.signature(s.route_request).parameters) == 4
# reconfigure (routing_logic.py:500-512) drops Session, keeps the prefix trie
s2 = rl.reconfigure_routing_logic(rl.RoutingLogic.SESSION_BASED, session_key="x-session")
assert s2 is not s and s2.session_key == "x-session"
assert gpu.PrefixAwareRouter() is r and FakeTrie.made == 1
r2 = rl.reconfigure_routing_logic(rl.RoutingLogic.PREFIXAWARE)
assert r2 is r and rl.get_routing_logic() is r                        # Session was dropped
# errors (routing_logic.py:179-180, 412)
rl.SingletonABCMeta._instances.pop(gpu.SessionRouter, None)
try:
    gpu.SessionRouter(None); raise SystemExit("SessionRouter(None) must raise")
except ValueError:
    pass
try:
    asyncio.run(r.route_request([], {}, {}, None, {})); raise SystemExit("KeyError expected")
except KeyError:
    pass
# session routing end to end on the stand-in ring, with the reference's fallback rule
class Ep:
    def __init__(self, url): self.url = url
class Req:
    def __init__(self, h): self.headers = h
class St:
    def __init__(self, q): self.qps = q
s3 = rl.initialize_routing_logic(rl.RoutingLogic.SESSION_BASED, session_key="sid")
eps = [Ep("http://a"), Ep("http://b")]
assert s3.route_request(eps, {}, {"http://a": St(3.0), "http://b": St(1.0)}, Req({})) == "http://b"
assert s3.route_request(eps, {}, {}, Req({"sid": "k"})) in ("http://a", "http://b")
assert sorted(s3.hash_ring.get_nodes()) == ["http://a", "http://b"]
# the other routing logics still come from the reference
rr = rl.initialize_routing_logic(rl.RoutingLogic.ROUND_ROBIN)
assert type(rr).__module__ == "vllm_router.routers.routing_logic"
print("DROPIN-OK")
'''


@pytest.mark.skipif(not os.path.isdir(REF_SRC), reason="reference tree not present")
def test_gpu_routers_drop_into_the_reference_registry():
    pytest.importorskip("xxhash")
    out = subprocess.run([sys.executable, "-c", textwrap.dedent(SCRIPT), ROOT, REF_SRC],
                         capture_output=True, text=True, timeout=300)
    assert out.returncode == 0 and "DROPIN-OK" in out.stdout, out.stdout[-2000:] + out.stderr[-4000:]
