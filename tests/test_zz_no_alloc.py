"""b200r_xxh64_chunks_noalloc: chunk hashing for callers that cannot allocate.

Some embedders of libb200router.so never get to call cudaMalloc: a routing
plugin inside a serving process whose engine has taken nearly all of HBM, or a
request-path thread for which the device-wide stall of cudaMalloc/cudaFree is
not acceptable.  They call this entry point, whose work space is a fixed 4 MiB
area inside the library instead of a caller-provided scratch buffer.

The test drives the entry point through the C ABI exactly as such a caller
would (inputs and outputs only) and requires bit-identical results -- these are
integer hashes, so the tolerance is zero -- to the regular b200r_xxh64_chunks on
two calls in a row: the second call reuses the area with whatever the first one
left behind.  The hashes are also held to the CPU oracle (an independent
implementation), and the device-wide free-memory figure must not drop across a
call -- that also catches a cudaMalloc made inside the library, which torch's own
allocator statistics would not see.  The file sorts last on purpose (pytest collects it after every
other test file), and it must pass on its own in a fresh process as well.
"""
import ctypes as C

import numpy as np
import pytest
import torch

from oracle import coracle
from b200_router import _lib, traces
from b200_router.packing import pack_arrays, pack_prompts

pytestmark = pytest.mark.gpu


def _vp(t):
    return None if t is None else C.c_void_p(t.data_ptr())


def _stream():
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


def _hash_regular(p, chunk=128):
    L = _lib.lib()
    dev = p.bytes.device
    base = torch.empty(p.n + 1, dtype=torch.int32, device=dev)
    hashes = torch.zeros(max(p.max_chunks, 1), dtype=torch.int64, device=dev)
    need = L.b200r_xxh64_chunks_scratch_bytes(p.n, p.max_chunks)
    scratch = torch.empty(need, dtype=torch.uint8, device=dev)
    flags = _lib.F_ALL_ASCII if p.all_ascii else 0
    _lib.check(L.b200r_xxh64_chunks(_vp(p.bytes), _vp(p.starts), _vp(p.lens), _vp(p.nchars), p.n,
                                    chunk, flags, _vp(base), _vp(hashes), p.max_chunks,
                                    _vp(scratch), need, _stream()))
    torch.cuda.synchronize()
    return base.cpu().numpy(), hashes.cpu().numpy()


def _hash_noalloc(p, chunk=128, check_free=False):
    """Inputs and outputs only -- no scratch argument, nothing allocated by the call.

    check_free: compare the DEVICE-WIDE free memory (cudaMemGetInfo) before and after the
    call.  Only meaningful once an earlier call has loaded the library's kernels and looked
    up the built-in area (the CUDA runtime loads a module lazily on first use)."""
    L = _lib.lib()
    dev = p.bytes.device
    base = torch.empty(p.n + 1, dtype=torch.int32, device=dev)
    hashes = torch.zeros(max(p.max_chunks, 1), dtype=torch.int64, device=dev)
    flags = _lib.F_ALL_ASCII if p.all_ascii else 0
    torch.cuda.synchronize()
    before = torch.cuda.memory_allocated()
    free_before = torch.cuda.mem_get_info(dev)[0]
    _lib.check(L.b200r_xxh64_chunks_noalloc(_vp(p.bytes), _vp(p.starts), _vp(p.lens),
                                            _vp(p.nchars), p.n, chunk, flags, _vp(base),
                                            _vp(hashes), p.max_chunks, _stream()))
    torch.cuda.synchronize()
    free_after = torch.cuda.mem_get_info(dev)[0]
    assert torch.cuda.memory_allocated() == before
    if check_free:
        assert free_after >= free_before, "device memory was allocated inside the call"
    return base.cpu().numpy(), hashes.cpu().numpy()


def _same_as_oracle(base, hs, prompts, chunk=128):
    """Every chunk hash equals the CPU oracle's (oracle/oracle.c, pinned to python-xxhash)."""
    hs = hs.view(np.uint64)
    for i, s in prompts:
        want = coracle.xxh64_chunks(s.encode("utf-8"), chunk)
        assert base[i + 1] - base[i] == len(want), i
        assert (hs[base[i]:base[i + 1]] == want).all(), i


def _same(a, b):
    total = int(a[0][-1])
    assert total == int(b[0][-1])
    assert (a[0] == b[0]).all()
    assert (a[1][:total] == b[1][:total]).all()


def test_noalloc_matches_regular_entry_point_twice_in_a_row():
    dev = torch.device("cuda", torch.cuda.current_device())
    # call 1: a large ASCII batch (C3-shaped), fills much of the tile map
    tr = traces.SharedPrefixTrace(n_total=20000, batch=20000, n_sys=16, sys_chars=2048,
                                  suffix_lo=256, suffix_hi=1024, seed=11)
    big = pack_arrays(*tr.batch_arrays(0), pin=False).to(dev)
    first = _hash_noalloc(big)
    _same(first, _hash_regular(big))
    buf, starts, lens = tr.batch_arrays(0)
    _same_as_oracle(first[0], first[1],
                    [(i, buf[int(starts[i]):int(starts[i]) + int(lens[i])].tobytes().decode("ascii"))
                     for i in range(0, 20000, 997)])
    # call 2, straight after: smaller, mixed ASCII / multi-byte UTF-8, ragged and empty
    # prompts -- uses the code-point counts and chunk offsets parts of the area as well
    rng = np.random.default_rng(5)
    alphabet = list("abc XYZ09") + ["é", "日", "🙂"]
    prompts = ["".join(rng.choice(alphabet, size=n)) for n in (0, 1, 127, 128, 129, 300, 1000, 5000)]
    prompts += ["", "x" * 4096, "é" * 128]
    small = pack_prompts(prompts, pin=False).to(dev)
    second = _hash_noalloc(small, check_free=True)
    _same(second, _hash_regular(small))
    _same_as_oracle(second[0], second[1], list(enumerate(prompts)))
    # and the first batch again: same answer as before, and no device memory taken
    _same(_hash_noalloc(big, check_free=True), first)


def test_noalloc_rejects_batches_beyond_the_built_in_area():
    dev = torch.device("cuda", torch.cuda.current_device())
    p = pack_prompts(["abc"], pin=False).to(dev)
    base = torch.empty(2, dtype=torch.int32, device=dev)
    hashes = torch.empty(1, dtype=torch.int64, device=dev)
    too_many = 4 * 1024 * 1024          # chunk capacity whose offsets alone need 16 MiB
    rc = _lib.lib().b200r_xxh64_chunks_noalloc(_vp(p.bytes), _vp(p.starts), _vp(p.lens),
                                               _vp(p.lens), 1, 128, 0, _vp(base), _vp(hashes),
                                               too_many, _stream())
    assert rc == -4                     # B200R_E_SCRATCH
    msg = _lib.lib().b200r_last_error().decode()
    assert "built-in area" in msg and str(_lib.NOALLOC_SCRATCH_BYTES) in msg
