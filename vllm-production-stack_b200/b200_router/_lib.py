"""ctypes binding of libb200router.so (include/b200_router.h).

The library is built in-tree by vllm-production-stack_b200/build.py and is the
only compute path: if it is missing, or no CUDA device is present, the routing
calls raise -- there is no CPU fallback.
"""
import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libb200router.so")

OK = 0
F_ALL_ASCII = 1
F_PREP_ONLY = 2
F_PREP_DONE = 4
S_TABLE_FULL = 1
S_CHUNK_CAP = 2
S_PEER_TIMEOUT = 4
S_INTERNAL = 8
NOALLOC_SCRATCH_BYTES = 4 * 1024 * 1024

_vp, _i64, _i32, _u32, _u64, _sz = C.c_void_p, C.c_int64, C.c_int32, C.c_uint32, C.c_uint64, C.c_size_t

# name -> (restype, argtypes); the symbol list is also what tests check against the header
SIGNATURES = {
    "b200r_abi_version": (C.c_int, []),
    "b200r_last_error": (C.c_char_p, []),
    "b200r_xxh64_chunks_scratch_bytes": (_sz, [_i64, _i64]),
    "b200r_xxh64_chunks": (C.c_int, [_vp, _vp, _vp, _vp, _i64, _i32, _u32, _vp, _vp, _i64, _vp, _sz, _vp]),
    "b200r_xxh64_chunks_noalloc": (C.c_int, [_vp, _vp, _vp, _vp, _i64, _i32, _u32, _vp, _vp, _i64, _vp]),
    "b200r_trie_create": (C.c_int, [_i32, _i32, C.POINTER(_vp)]),
    "b200r_trie_destroy": (C.c_int, [_vp]),
    "b200r_trie_clear": (C.c_int, [_vp, _vp]),
    "b200r_trie_status": (C.c_int, [_vp, _vp, C.POINTER(_u32), C.POINTER(_i64), C.POINTER(_u64),
                                    C.POINTER(_i64)]),
    "b200r_trie_status_async": (C.c_int, [_vp, _vp, _vp]),
    "b200r_trie_retire_endpoints": (C.c_int, [_vp, _u64, _vp]),
    "b200r_trie_set_epoch": (C.c_int, [_vp, _u32]),
    "b200r_trie_lookup_batch": (C.c_int, [_vp, _vp, _vp, _i64, _u64, _vp, _vp, _vp]),
    "b200r_trie_insert_batch": (C.c_int, [_vp, _vp, _vp, _i64, _vp, _vp]),
    "b200r_prefix_route_scratch_bytes": (_sz, [_i64, _i64, _u32]),
    "b200r_prefix_route_batch": (C.c_int, [_vp, _vp, _vp, _vp, _vp, _i64, _i64, _u32, _u64, _u64, _u64,
                                           _vp, _vp, _vp, _vp, _sz, _vp]),
    "b200r_prefix_hash_stage": (C.c_int, [_vp, _vp, _vp, _vp, _vp, _i64, _i64, _u32, _vp, _sz, _vp]),
    "b200r_prefix_route_stage": (C.c_int, [_vp, _i64, _i64, _u32, _u64, _u64, _u64, _vp, _vp, _vp, _vp,
                                           _sz, _vp]),
    "b200r_prefix_route_hashes": (C.c_int, [_vp, _vp, _vp, _i64, _i64, _u64, _u64, _u64, _vp, _vp, _vp, _vp,
                                            _sz, _vp]),
    "b200r_prefix_route_one": (C.c_int, [_vp, C.c_char_p, _i64, _u64, _u64, _u64, C.POINTER(C.c_int32),
                                         C.POINTER(C.c_int32), C.POINTER(C.c_uint64),
                                         C.POINTER(C.c_uint32), _vp]),
    "b200r_pipe_create": (C.c_int, [_vp, _i32, C.POINTER(_vp)]),
    "b200r_pipe_destroy": (C.c_int, [_vp]),
    "b200r_pipe_prep_stream": (_vp, [_vp]),
    "b200r_pipe_begin": (C.c_int, [_vp]),
    "b200r_pipe_copy_h2d": (C.c_int, [_vp, _vp, _vp, _sz]),
    "b200r_pipe_route": (C.c_int, [_vp, _vp, _vp, _vp, _vp, _i64, _i64, _u32, _u64, _u64, _u64, _vp, _vp, _vp,
                                   _vp, _sz, _i32, _vp]),
    "b200r_pick": (_u32, [_u64, _u64, _u32]),
    "b200r_trie_dump": (C.c_int, [_vp, _vp, _i64, C.POINTER(_i64), C.POINTER(_u64), _vp]),
    "b200r_trie_restore": (C.c_int, [_vp, _vp, _i64, _u64, _vp]),
    "b200r_part_layout": (C.c_int, [_i32, _i64, _i64, C.POINTER(_sz)]),
    "b200r_part_send": (C.c_int, [_vp, _vp, _vp, _vp, _vp, _i64, _i64, _u32, _vp, _i32, _i32, _i64, _i64,
                                  _u32, _vp, _vp, _sz, _vp]),
    "b200r_part_route": (C.c_int, [_vp, _u64, _u64, _vp, _vp, _i32, _i32, _i64, _i64, _u32, _vp, _vp]),
    "b200r_part_finish": (C.c_int, [_vp, _i64, _vp, _i32, _i32, _i64, _i64, _u32, _vp, _vp, _vp, _vp, _vp]),
    "b200r_part_begin": (C.c_int, [_vp, _u32]),
    "b200r_part_copy_h2d": (C.c_int, [_vp, _vp, _vp, _sz]),
    "b200r_part_step": (C.c_int, [_vp, _vp, _vp, _vp, _vp, _i64, _i64, _u32, _u64, _u64, _vp, _vp, _i32, _i32,
                                  _i64, _i64, _u32, _vp, _vp, _sz, _vp, _vp, _vp, _i32, _vp]),
    "b200r_set_part_staged": (C.c_int, [C.c_int]),
    "b200r_set_hash_mode": (C.c_int, [C.c_int]),
    "b200r_set_claim_mode": (C.c_int, [C.c_int]),
    "b200r_prof_enable": (C.c_int, [C.c_int]),
    "b200r_prof_read": (C.c_int, [C.POINTER(C.c_float), C.POINTER(C.c_longlong), C.c_int]),
    "b200r_launch_count": (C.c_longlong, []),
    "b200r_md5_batch": (C.c_int, [_vp, _vp, _vp, _i64, _vp, _vp]),
    "b200r_ring_create": (C.c_int, [C.POINTER(_vp)]),
    "b200r_ring_destroy": (C.c_int, [_vp]),
    "b200r_ring_build": (C.c_int, [_vp, _vp, _vp, _vp, _i32, _vp]),
    "b200r_ring_lookup_batch": (C.c_int, [_vp, _vp, _vp, _vp, _i64, _vp, _vp]),
}

_lib = None


class B200RouterError(RuntimeError):
    pass


def lib():
    """Load the shared library (raises if it has not been built)."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise B200RouterError(
                "libb200router.so not found at %s: run `python vllm-production-stack_b200/build.py` "
                "(needs nvcc); there is no CPU fallback" % LIB_PATH)
        L = C.CDLL(LIB_PATH)
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(L, name)
            fn.restype = res
            fn.argtypes = args
        if L.b200r_abi_version() != 2:
            raise B200RouterError("libb200router.so ABI mismatch")
        _lib = L
    return _lib


def check(rc: int):
    if rc != OK:
        msg = lib().b200r_last_error().decode("utf-8", "replace")
        raise B200RouterError("libb200router error %d: %s" % (rc, msg))


def require_cuda():
    import torch
    if not torch.cuda.is_available():
        raise B200RouterError("no CUDA device: the B200 router has no CPU fallback")
    return torch


PROF_SLOTS = ("prep", "utf8", "hash", "claim", "select", "commit", "lookup", "insert",
              "classify", "push", "wait", "apply", "resolve")


def prof_enable(on: bool):
    check(lib().b200r_prof_enable(1 if on else 0))


def prof_read():
    """{kernel class: (ms, launches)} since the last read (synchronises)."""
    ms = (C.c_float * len(PROF_SLOTS))()
    cnt = (C.c_longlong * len(PROF_SLOTS))()
    check(lib().b200r_prof_read(ms, cnt, len(PROF_SLOTS)))
    return {k: (float(ms[i]), int(cnt[i])) for i, k in enumerate(PROF_SLOTS)}


def launch_count() -> int:
    return int(lib().b200r_launch_count())
