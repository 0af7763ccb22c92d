"""Shared helpers of the parity tests (GPU path through the C ABI vs the oracle)."""
import numpy as np
import torch

from b200_router import traces
from b200_router.packing import pack_arrays, pack_prompts


def bits(m):
    m = int(m) & ((1 << 64) - 1)
    return [i for i in range(64) if (m >> i) & 1]


def u64(t: torch.Tensor) -> np.ndarray:
    return t.cpu().numpy().view(np.uint64)


def trie_as_sorted_list(dump):
    """{path tuple: mask} -> the sorted list layout of tests/golden/prefix_traces.json"""
    return sorted([["%016x" % h for h in path], bits(m)] for path, m in dump.items())


def tight_pack(prompts):
    """CSR-tight packing (unaligned starts) to exercise the generic path."""
    blobs = [p.encode("utf-8") for p in prompts]
    lens = np.array([len(b) for b in blobs], dtype=np.int32)
    nchars = np.array([len(p) for p in prompts], dtype=np.int32)
    starts = np.zeros(len(blobs), dtype=np.int64)
    if len(blobs) > 1:
        starts[1:] = np.cumsum(lens[:-1])
    buf = np.frombuffer(b"".join(blobs) + b"\0" * 16, dtype=np.uint8).copy()
    return pack_arrays(buf, starts, lens, nchars, pin=False)
