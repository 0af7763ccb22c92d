"""Packed prompt layout shared by every batch entry point.

"PyTorch tensors hold packed prompt bytes plus offsets" (north_star): request i
is bytes[starts[i] : starts[i] + lens[i]] (UTF-8).  The packer aligns every
start to 16 bytes so that the hashing kernel's cp.async staging applies; nchars
carries Python's len(str) (code points) because the reference chunks the *str*
(hashtrie.py:55) and hashes the UTF-8 encoding of each slice.
"""
import ctypes as C
import os
import threading
from dataclasses import dataclass
from typing import List, Optional, Sequence

import numpy as np
import torch

ALIGN = 16
_PADS = [b"\0" * k for k in range(ALIGN)]


@dataclass
class PackedPrompts:
    bytes: torch.Tensor              # uint8 [total]
    starts: torch.Tensor             # int64 [n]
    lens: torch.Tensor               # int32 [n]  (bytes)
    nchars: Optional[torch.Tensor]   # int32 [n]  (code points); None == same as lens
    n: int
    max_chunks: int                  # sum ceil(nchars_i / chunk_chars)
    all_ascii: bool
    payload_bytes: int               # sum(lens)

    @property
    def device(self):
        return self.bytes.device

    def to(self, device, non_blocking: bool = True) -> "PackedPrompts":
        mv = lambda t: None if t is None else t.to(device, non_blocking=non_blocking)
        return PackedPrompts(mv(self.bytes), mv(self.starts), mv(self.lens), mv(self.nchars),
                             self.n, self.max_chunks, self.all_ascii, self.payload_bytes)

    def h2d_bytes(self) -> int:
        tot = self.bytes.numel() + self.starts.numel() * 8 + self.lens.numel() * 4
        if self.nchars is not None:
            tot += self.nchars.numel() * 4
        return tot


def _maybe_pin(t: torch.Tensor, pin: bool) -> torch.Tensor:
    if pin and torch.cuda.is_available():
        return t.pin_memory()
    return t


def _as_str(p, sep: str) -> str:
    return p if isinstance(p, str) else sep.join(p)


def _pack_prompts_py(prompts, chunk_chars, pin, by_bytes, sep) -> PackedPrompts:
    """Pure-Python packer (reference for the native one; used when libb200pack.so is absent)."""
    n = len(prompts)
    strs = [_as_str(p, sep) for p in prompts]
    blobs: List[bytes] = [p.encode("utf-8") for p in strs]
    lens = np.fromiter((len(b) for b in blobs), dtype=np.int32, count=n)
    nchars = lens.copy() if by_bytes else np.fromiter((len(p) for p in strs), dtype=np.int32, count=n)
    pitch = (lens.astype(np.int64) + ALIGN - 1) // ALIGN * ALIGN
    starts = np.zeros(n, dtype=np.int64)
    if n > 1:
        np.cumsum(pitch[:-1], out=starts[1:])
    pieces = []
    for b in blobs:
        pieces.append(b)
        pieces.append(_PADS[-len(b) % ALIGN])
    blob = b"".join(pieces)
    buf = np.frombuffer(blob, dtype=np.uint8) if blob else np.zeros(ALIGN, dtype=np.uint8)
    all_ascii = bool((lens == nchars).all())
    max_chunks = int(((nchars.astype(np.int64) + chunk_chars - 1) // chunk_chars).sum())
    return PackedPrompts(
        _maybe_pin(torch.from_numpy(buf.copy()), pin), _maybe_pin(torch.from_numpy(starts), pin),
        _maybe_pin(torch.from_numpy(lens), pin),
        None if all_ascii else _maybe_pin(torch.from_numpy(nchars), pin),
        n, max_chunks, all_ascii, int(lens.sum()))


_PACK_LIB = None
PACK_LIB_PATH = os.path.join(os.path.dirname(os.path.abspath(__file__)), "libb200pack.so")


def pack_lib():
    """libb200pack.so (csrc/packer.cpp, include/b200_pack.h) or None when it is not built.
    PyDLL: the calls take borrowed PyObject references and keep the GIL."""
    global _PACK_LIB
    if _PACK_LIB is None:
        if not os.path.exists(PACK_LIB_PATH):
            _PACK_LIB = False
        else:
            L = C.PyDLL(PACK_LIB_PATH)
            L.b200p_last_error.restype = C.c_char_p
            L.b200p_measure.restype = C.c_int
            L.b200p_measure.argtypes = [C.py_object, C.py_object, C.c_int64, C.c_int, C.POINTER(C.c_int64)]
            L.b200p_fill.restype = C.c_int
            L.b200p_fill.argtypes = [C.py_object, C.py_object, C.c_void_p, C.c_int64, C.c_void_p,
                                     C.c_void_p, C.c_void_p, C.c_int, C.c_int]
            _PACK_LIB = L
    return _PACK_LIB or None


class PackArena:
    """Reusable (pinned) host buffers for pack_prompts: allocation of pinned memory costs far
    more than the packing itself.  `slots` buffers are handed out in rotation, so a batch may
    still be in flight to the device while the next `slots - 1` are being packed."""

    def __init__(self, slots: int = 2, pin: bool = True):
        self.pin = pin and torch.cuda.is_available()
        self._slots = [dict() for _ in range(max(slots, 1))]
        self._i = 0
        self._lock = threading.Lock()

    def _get(self, slot, name, numel, dtype):
        t = slot.get(name)
        if t is None or t.numel() < numel:
            t = slot[name] = torch.empty(int(numel * 1.25) + 64, dtype=dtype, pin_memory=self.pin)
        return t[:numel]

    def take(self, nbytes: int, n: int):
        with self._lock:
            slot = self._slots[self._i % len(self._slots)]
            self._i += 1
        return (self._get(slot, "bytes", nbytes, torch.uint8), self._get(slot, "starts", n, torch.int64),
                self._get(slot, "lens", n, torch.int32), self._get(slot, "nchars", n, torch.int32))


class PackPipeline:
    """Packs batches ahead of their use on worker threads.  The copy phase of the native packer
    runs without the GIL, so two batches in flight roughly double the packing rate; results
    come back in submission order."""

    def __init__(self, arena: Optional["PackArena"] = None, workers: int = 2, **pack_kwargs):
        from concurrent.futures import ThreadPoolExecutor
        self.arena = arena if arena is not None else PackArena(slots=workers + 2)
        self.kwargs = pack_kwargs
        self._pool = ThreadPoolExecutor(max_workers=max(workers, 1))
        self._queue = []

    def submit(self, prompts):
        # every batch in the queue, and every result the caller still uses, holds one arena slot
        if len(self._queue) >= len(self.arena._slots) - 1:
            raise RuntimeError("PackPipeline: %d batches already queued for an arena of %d slots"
                               % (len(self._queue), len(self.arena._slots)))
        self._queue.append(self._pool.submit(pack_prompts, prompts, arena=self.arena, **self.kwargs))

    def next(self) -> "PackedPrompts":
        return self._queue.pop(0).result()

    def close(self):
        self._pool.shutdown(wait=True)


def pack_prompts(prompts: Sequence, chunk_chars: int = 128, pin: bool = True, chunk_by: str = "chars",
                 sep: str = "\n", arena: Optional[PackArena] = None, threads: int = 0) -> PackedPrompts:
    """Pack prompts on the host (pinned when a GPU is present).

    A prompt is a str, or a list/tuple of str parts standing for sep.join(parts) -- the chat
    messages' contents of routing_logic.py:389-407 -- which is packed without building the joined
    string.  chunk_by="chars" chunks by code points like the reference router (hashtrie.py:55);
    "bytes" chunks the UTF-8 bytes like the gateway picker (prefix_aware_picker.go:154-165).
    The work is done by the native packer (csrc/packer.cpp) when it is built; `threads` = copy
    threads for large batches (0 = automatic, up to 8)."""
    if chunk_by not in ("chars", "bytes"):
        raise ValueError("chunk_by must be 'chars' or 'bytes'")
    by_bytes = chunk_by == "bytes"
    if not isinstance(prompts, (list, tuple)):
        prompts = list(prompts)
    L = pack_lib()
    if L is None:
        return _pack_prompts_py(prompts, chunk_chars, pin, by_bytes, sep)
    n = len(prompts)
    m = (C.c_int64 * 4)()
    if L.b200p_measure(prompts, sep, chunk_chars, int(by_bytes), m):
        raise TypeError(L.b200p_last_error().decode())
    total, payload, max_chunks, ascii_ = int(m[0]), int(m[1]), int(m[2]), bool(m[3])
    if arena is not None:
        buf, starts, lens, nchars = arena.take(total, n)
    else:
        pinned = pin and torch.cuda.is_available()
        buf = torch.empty(total, dtype=torch.uint8, pin_memory=pinned)
        starts = torch.empty(n, dtype=torch.int64, pin_memory=pinned)
        lens = torch.empty(n, dtype=torch.int32, pin_memory=pinned)
        nchars = torch.empty(n, dtype=torch.int32, pin_memory=pinned)
    if L.b200p_fill(prompts, sep, buf.data_ptr(), total, starts.data_ptr(), lens.data_ptr(),
                    nchars.data_ptr(), int(by_bytes), threads):
        raise TypeError(L.b200p_last_error().decode())
    same = ascii_ or by_bytes
    return PackedPrompts(buf, starts, lens, None if same else nchars, n, max_chunks, same, payload)


def pack_arrays(buf: np.ndarray, starts: np.ndarray, lens: np.ndarray, nchars=None,
                chunk_chars: int = 128, pin: bool = True) -> PackedPrompts:
    """Wrap already-packed numpy arrays (trace generators)."""
    lens = np.ascontiguousarray(lens, dtype=np.int32)
    starts = np.ascontiguousarray(starts, dtype=np.int64)
    buf = np.ascontiguousarray(buf, dtype=np.uint8)
    if buf.size == 0:
        buf = np.zeros(ALIGN, dtype=np.uint8)
    nc = lens if nchars is None else np.ascontiguousarray(nchars, dtype=np.int32)
    all_ascii = nchars is None or bool((nc == lens).all())
    max_chunks = int(((nc.astype(np.int64) + chunk_chars - 1) // chunk_chars).sum())
    return PackedPrompts(
        _maybe_pin(torch.from_numpy(buf), pin), _maybe_pin(torch.from_numpy(starts), pin),
        _maybe_pin(torch.from_numpy(lens), pin),
        None if all_ascii else _maybe_pin(torch.from_numpy(nc), pin),
        len(lens), max_chunks, all_ascii, int(lens.sum()))
