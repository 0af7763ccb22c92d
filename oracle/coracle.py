"""ctypes front-end of oracle.c (liboracle.so).  TEST INFRASTRUCTURE ONLY."""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "liboracle.so")
_lib = None

u8p = C.POINTER(C.c_uint8)
i64p = C.POINTER(C.c_int64)
i32p = C.POINTER(C.c_int32)
u64p = C.POINTER(C.c_uint64)
u32p = C.POINTER(C.c_uint32)


def build(force: bool = False) -> str:
    src = os.path.join(_HERE, "oracle.c")
    if force or not os.path.exists(_SO) or os.path.getmtime(_SO) < os.path.getmtime(src):
        subprocess.check_call(["make", "-C", _HERE, "-B", "liboracle.so"],
                              stdout=subprocess.DEVNULL)
    return _SO


def lib():
    global _lib
    if _lib is None:
        build()                      # no-op unless oracle.c is newer than the library
        L = C.CDLL(_SO)
        L.orc_xxh64.restype = C.c_uint64
        L.orc_xxh64.argtypes = [C.c_void_p, C.c_int64, C.c_uint64]
        L.orc_xxh64_chunks.restype = C.c_int64
        L.orc_xxh64_chunks.argtypes = [C.c_void_p, C.c_int64, C.c_int, C.c_void_p, C.c_int64]
        L.orc_trie_new.restype = C.c_void_p
        L.orc_trie_free.argtypes = [C.c_void_p]
        L.orc_trie_lookup.restype = C.c_int
        L.orc_trie_lookup.argtypes = [C.c_void_p, C.c_void_p, C.c_int64, C.c_int, C.c_uint64, u64p]
        L.orc_trie_insert.argtypes = [C.c_void_p, C.c_void_p, C.c_int64, C.c_int, C.c_int]
        L.orc_pick.restype = C.c_uint32
        L.orc_pick.argtypes = [C.c_uint64, C.c_uint64, C.c_uint32]
        L.orc_route.restype = C.c_int
        L.orc_route.argtypes = [C.c_void_p, C.c_void_p, C.c_int64, C.c_int, C.c_uint64,
                                C.c_uint64, C.c_uint64, i32p, u64p]
        L.orc_route_batch.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64,
                                      C.c_int, C.c_uint64, C.c_uint64, C.c_uint64,
                                      C.c_void_p, C.c_void_p, C.c_void_p]
        L.orc_lookup_batch.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64,
                                       C.c_int, C.c_uint64, C.c_void_p, C.c_void_p]
        L.orc_insert_batch.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64,
                                       C.c_int, C.c_void_p]
        L.orc_trie_num_nodes.restype = C.c_int64
        L.orc_trie_num_nodes.argtypes = [C.c_void_p]
        L.orc_trie_dump.argtypes = [C.c_void_p] + [C.c_void_p] * 4
        L.orc_md5.argtypes = [C.c_void_p, C.c_int64, C.c_void_p]
        L.orc_md5_u128_batch.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64,
                                         C.c_void_p, C.c_void_p]
        L.orc_ring_lookup_batch.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64,
                                            C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64,
                                            C.c_void_p]
        _lib = L
    return _lib


def _ptr(a):
    return a.ctypes.data_as(C.c_void_p)


def xxh64(data: bytes, seed: int = 0) -> int:
    buf = (C.c_uint8 * max(1, len(data))).from_buffer_copy(data or b"\0")
    return lib().orc_xxh64(buf, len(data), seed)


def xxh64_chunks(data: bytes, chunk_chars: int = 128, by_bytes: bool = False):
    cap = len(data) // chunk_chars + 2
    out = np.zeros(cap, dtype=np.uint64)
    buf = np.frombuffer(data or b"\0", dtype=np.uint8)
    lib().orc_set_chunk_by_bytes(int(by_bytes))
    n = lib().orc_xxh64_chunks(_ptr(buf), len(data), chunk_chars, _ptr(out), cap)
    assert n <= cap
    return out[:n]


def pick(seed: int, seq: int, n: int) -> int:
    return lib().orc_pick(seed, seq, n)


def md5(data: bytes) -> bytes:
    out = (C.c_uint8 * 16)()
    buf = (C.c_uint8 * max(1, len(data))).from_buffer_copy(data or b"\0")
    lib().orc_md5(buf, len(data), out)
    return bytes(out)


class Packed:
    """Host-side packed prompts: bytes + per-request start/len (numpy)."""

    def __init__(self, blobs):
        lens = np.array([len(b) for b in blobs], dtype=np.int32)
        starts = np.zeros(len(blobs), dtype=np.int64)
        if len(blobs):
            starts[1:] = np.cumsum(lens[:-1], dtype=np.int64)
        self.bytes = np.frombuffer(b"".join(blobs) or b"\0", dtype=np.uint8)
        self.starts, self.lens, self.n = starts, lens, len(blobs)


class TrieOracle:
    """Sequential pointer trie with endpoint bitmasks (oracle.c)."""

    def __init__(self, chunk_chars: int = 128, by_bytes: bool = False):
        """by_bytes: chunk the UTF-8 bytes (gateway picker) instead of the code points (router)."""
        self.h = lib().orc_trie_new()
        self.chunk_chars = chunk_chars
        self.by_bytes = by_bytes

    def __del__(self):
        try:
            if getattr(self, "h", None):
                lib().orc_trie_free(self.h)
                self.h = None
        except Exception:      # interpreter shutdown
            pass

    def route_batch(self, byts, starts, lens, avail, seed, seq0):
        n = len(starts)
        ep = np.empty(n, dtype=np.int32)
        depth = np.empty(n, dtype=np.int32)
        matched = np.empty(n, dtype=np.uint64)
        byts = np.ascontiguousarray(byts, dtype=np.uint8)
        starts = np.ascontiguousarray(starts, dtype=np.int64)
        lens = np.ascontiguousarray(lens, dtype=np.int32)
        lib().orc_set_chunk_by_bytes(int(self.by_bytes))
        lib().orc_route_batch(self.h, _ptr(byts), _ptr(starts), _ptr(lens), n, self.chunk_chars,
                              avail, seed, seq0, _ptr(ep), _ptr(depth), _ptr(matched))
        return ep, depth, matched

    def lookup_batch(self, byts, starts, lens, avail):
        n = len(starts)
        depth = np.empty(n, dtype=np.int32)
        matched = np.empty(n, dtype=np.uint64)
        byts = np.ascontiguousarray(byts, dtype=np.uint8)
        starts = np.ascontiguousarray(starts, dtype=np.int64)
        lens = np.ascontiguousarray(lens, dtype=np.int32)
        lib().orc_set_chunk_by_bytes(int(self.by_bytes))
        lib().orc_lookup_batch(self.h, _ptr(byts), _ptr(starts), _ptr(lens), n, self.chunk_chars,
                               avail, _ptr(depth), _ptr(matched))
        return depth, matched

    def insert_batch(self, byts, starts, lens, endpoint):
        byts = np.ascontiguousarray(byts, dtype=np.uint8)
        starts = np.ascontiguousarray(starts, dtype=np.int64)
        lens = np.ascontiguousarray(lens, dtype=np.int32)
        endpoint = np.ascontiguousarray(endpoint, dtype=np.int32)
        lib().orc_set_chunk_by_bytes(int(self.by_bytes))
        lib().orc_insert_batch(self.h, _ptr(byts), _ptr(starts), _ptr(lens), len(starts),
                               self.chunk_chars, _ptr(endpoint))

    def num_nodes(self) -> int:
        return lib().orc_trie_num_nodes(self.h)

    def dump_arrays(self):
        n = self.num_nodes()
        parent = np.empty(n, dtype=np.uint32)
        h = np.empty(n, dtype=np.uint64)
        depth = np.empty(n, dtype=np.uint32)
        mask = np.empty(n, dtype=np.uint64)
        lib().orc_trie_dump(self.h, _ptr(parent), _ptr(h), _ptr(depth), _ptr(mask))
        return parent, h, depth, mask

    def dump(self):
        """{tuple(chunk hashes from the root): mask}; () is the root."""
        parent, h, depth, mask = self.dump_arrays()
        paths = [()] * len(parent)
        out = {(): int(mask[0])}
        for i in range(1, len(parent)):        # parents are created before children
            paths[i] = paths[parent[i]] + (int(h[i]),)
            out[paths[i]] = int(mask[i])
        return out


def md5_u128_batch(byts, starts, lens):
    n = len(starts)
    hi = np.empty(n, dtype=np.uint64)
    lo = np.empty(n, dtype=np.uint64)
    lib().orc_md5_u128_batch(_ptr(np.ascontiguousarray(byts, dtype=np.uint8)),
                             _ptr(np.ascontiguousarray(starts, dtype=np.int64)),
                             _ptr(np.ascontiguousarray(lens, dtype=np.int32)), n, _ptr(hi), _ptr(lo))
    return hi, lo


def ring_lookup_batch(ring_hi, ring_lo, owner, byts, starts, lens):
    n = len(starts)
    out = np.empty(n, dtype=np.int32)
    ring_hi = np.ascontiguousarray(ring_hi, dtype=np.uint64)
    ring_lo = np.ascontiguousarray(ring_lo, dtype=np.uint64)
    owner = np.ascontiguousarray(owner, dtype=np.int32)
    lib().orc_ring_lookup_batch(_ptr(ring_hi), _ptr(ring_lo), _ptr(owner), len(owner),
                                _ptr(np.ascontiguousarray(byts, dtype=np.uint8)),
                                _ptr(np.ascontiguousarray(starts, dtype=np.int64)),
                                _ptr(np.ascontiguousarray(lens, dtype=np.int32)), n, _ptr(out))
    return out
