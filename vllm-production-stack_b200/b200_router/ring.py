"""GPU-resident consistent-hash ring with the API subset of uhashring.HashRing
that the reference router uses (routing_logic.py:92,99,103,217):
get_nodes / add_node / remove_node / get_node, plus a batched lookup.

Semantics restated from uhashring==2.3's default (non-ketama) ring: 160 points
per node at int(md5(f"{node}-{w}").hexdigest(), 16); lookup = bisect_right over
the sorted points, wrapping to the first.  The md5s -- ring points as well as
lookups -- are computed by the CUDA kernels in csrc/session.cu.
"""
import ctypes as C
from typing import Dict, List, Sequence

import numpy as np
import torch

from . import _lib
from .packing import PackedPrompts, pack_prompts
from .trie import _ptr, _stream

VNODES = 160


class GpuHashRing:
    def __init__(self, nodes: Sequence[str] = (), device=None):
        torch_ = _lib.require_cuda()
        self.device = torch_.device("cuda", torch_.cuda.current_device()) if device is None \
            else torch_.device(device)
        self._h = C.c_void_p()
        _lib.check(_lib.lib().b200r_ring_create(C.byref(self._h)))
        self._names: List[str] = []
        self._index: Dict[str, int] = {}      # node name -> stable owner index
        self._owner_name: List[str] = []      # owner index -> node name
        self._points: Dict[int, int] = {}     # 128-bit point -> owner index
        self._dirty = True
        for n in nodes:
            self.add_node(n)

    def close(self):
        if getattr(self, "_h", None) and self._h.value:
            _lib.lib().b200r_ring_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # --------------------------------------------------------------- md5 (GPU)
    def md5_ints(self, keys: Sequence[str]) -> List[int]:
        p = pack_prompts(list(keys), pin=False).to(self.device)
        out = self.md5_packed(p).cpu().numpy().view(np.uint64)
        return [(int(out[2 * i]) << 64) | int(out[2 * i + 1]) for i in range(len(keys))]

    def md5_packed(self, p: PackedPrompts) -> torch.Tensor:
        out = torch.empty(max(p.n, 1) * 2, dtype=torch.int64, device=self.device)
        with torch.cuda.device(self.device):
            _lib.check(_lib.lib().b200r_md5_batch(_ptr(p.bytes), _ptr(p.starts), _ptr(p.lens), p.n,
                                                  _ptr(out), _stream()))
        return out[:2 * p.n]

    # -------------------------------------------------------- uhashring subset
    def get_nodes(self) -> List[str]:
        return list(self._names)

    def add_node(self, name: str):
        if name in self._names:
            return
        if name not in self._index:
            self._index[name] = len(self._owner_name)
            self._owner_name.append(name)
        self._names.append(name)
        idx = self._index[name]
        for pt in self.md5_ints(["%s-%s" % (name, w) for w in range(VNODES)]):
            self._points[pt] = idx
        self._dirty = True

    def remove_node(self, name: str):
        if name not in self._names:
            raise KeyError("node '%s' not found, available nodes: %s" % (name, self._names))
        self._names.remove(name)
        for pt in self.md5_ints(["%s-%s" % (name, w) for w in range(VNODES)]):
            self._points.pop(pt, None)
        self._dirty = True

    def _upload(self):
        keys = sorted(self._points)
        hi = np.array([k >> 64 for k in keys], dtype=np.uint64)
        lo = np.array([k & ((1 << 64) - 1) for k in keys], dtype=np.uint64)
        owner = np.array([self._points[k] for k in keys], dtype=np.int32)
        with torch.cuda.device(self.device):
            _lib.check(_lib.lib().b200r_ring_build(
                self._h, hi.ctypes.data_as(C.c_void_p), lo.ctypes.data_as(C.c_void_p),
                owner.ctypes.data_as(C.c_void_p), len(keys), _stream()))
        self._dirty = False

    def lookup_packed(self, p: PackedPrompts) -> torch.Tensor:
        """Owner index (int32 device tensor) per key; -1 if the ring is empty."""
        if self._dirty:
            self._upload()
        if p.bytes.device != self.device:
            p = p.to(self.device)
        out = torch.empty(max(p.n, 1), dtype=torch.int32, device=self.device)
        with torch.cuda.device(self.device):
            _lib.check(_lib.lib().b200r_ring_lookup_batch(self._h, _ptr(p.bytes), _ptr(p.starts),
                                                          _ptr(p.lens), p.n, _ptr(out), _stream()))
        return out[:p.n]

    def get_node_batch(self, keys: Sequence[str]) -> List:
        idx = self.lookup_packed(pack_prompts([str(k) for k in keys], pin=False)).cpu().tolist()
        return [self._owner_name[i] if i >= 0 else None for i in idx]

    def get_node(self, key):
        if not self._points:
            return None
        return self.get_node_batch([key])[0]

    def owner_names(self) -> List[str]:
        return list(self._owner_name)
