"""Packed prompt layout shared by every batch entry point.

"PyTorch tensors hold packed prompt bytes plus offsets" (north_star): request i
is bytes[starts[i] : starts[i] + lens[i]] (UTF-8).  The packer aligns every
start to 16 bytes so that the hashing kernel's cp.async staging applies; nchars
carries Python's len(str) (code points) because the reference chunks the *str*
(hashtrie.py:55) and hashes the UTF-8 encoding of each slice.
"""
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


def pack_prompts(prompts: Sequence[str], chunk_chars: int = 128, pin: bool = True) -> PackedPrompts:
    """Pack python strings on the host (pinned when a GPU is present)."""
    n = len(prompts)
    blobs: List[bytes] = [p.encode("utf-8") for p in prompts]
    lens = np.fromiter((len(b) for b in blobs), dtype=np.int32, count=n)
    nchars = np.fromiter((len(p) for p in prompts), dtype=np.int32, count=n)
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
