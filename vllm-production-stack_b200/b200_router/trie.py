"""GPU-resident replacement of the reference HashTrie.

Mirrors src/vllm_router/prefix/hashtrie.py (HashTrie: _chunk_and_hash, insert,
longest_prefix_match) with batch methods over packed prompts; every method calls
the C ABI (include/b200_router.h) on the current CUDA stream.  No CPU fallback.
"""
import ctypes as C
from typing import Dict, Optional, Tuple

import numpy as np
import torch

from . import _lib
from .packing import PackedPrompts

M64 = (1 << 64) - 1


def _ptr(t: Optional[torch.Tensor]):
    return None if t is None else C.c_void_p(t.data_ptr())


def _stream():
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


def claim_kernel_name(n: int, chunks: int) -> str:
    """The kernel the automatic claim decomposition picks for a batch of n requests with
    `chunks` chunks in total (csrc/prefix.cu: claim_kind_for)."""
    return "k_claim_thread" if chunks <= n * 64 else "k_walk<claim>"


class GpuHashTrie:
    """Open-addressing table of trie nodes on one GPU.

    Node log + hash table of run heads (csrc/trie.cuh); capacity = 2**cap_log2 nodes, 64 bytes
    of device memory per node of capacity.  The reference never evicts
    (routing_logic.py:352) and neither does the log; `status()` reports the node count, a
    full log raises."""

    def __init__(self, cap_log2: int = 22, chunk_chars: int = 128, device=None):
        torch_ = _lib.require_cuda()
        self.device = torch_.device("cuda", torch_.cuda.current_device()) if device is None \
            else torch_.device(device)
        self.cap_log2 = cap_log2
        self.chunk_chars = chunk_chars
        self._h = C.c_void_p()
        with torch.cuda.device(self.device):
            _lib.check(_lib.lib().b200r_trie_create(cap_log2, chunk_chars, C.byref(self._h)))
        self._scratch = None

    def close(self):
        if getattr(self, "_h", None) and self._h.value:
            _lib.lib().b200r_trie_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # ------------------------------------------------------------------ utils
    def _scratch_for(self, nbytes: int) -> torch.Tensor:
        if self._scratch is None or self._scratch.numel() < nbytes:
            self._scratch = torch.empty(int(nbytes * 1.25) + 4096, dtype=torch.uint8,
                                        device=self.device)
        return self._scratch

    def reserve(self, n: int, max_chunks: int, all_ascii: bool = True):
        """Pre-size the scratch buffer (so the hot loop never allocates)."""
        flags = _lib.F_ALL_ASCII if all_ascii else 0
        self._scratch_for(_lib.lib().b200r_prefix_route_scratch_bytes(n, max_chunks, flags))

    def _dev(self, p: PackedPrompts) -> PackedPrompts:
        return p if p.bytes.device == self.device else p.to(self.device)

    def clear(self):
        with torch.cuda.device(self.device):
            _lib.check(_lib.lib().b200r_trie_clear(self._h, _stream()))

    def status(self, count_nodes: bool = True) -> Dict[str, int]:
        """Latched status bits, root mask, capacity; the node count is a table
        scan on the device and only done when asked for."""
        st, nn, rm, cap = C.c_uint32(), C.c_int64(-1), C.c_uint64(), C.c_int64()
        with torch.cuda.device(self.device):
            _lib.check(_lib.lib().b200r_trie_status(self._h, _stream(), C.byref(st),
                                                    C.byref(nn) if count_nodes else None,
                                                    C.byref(rm), C.byref(cap)))
        return {"status": st.value, "num_nodes": nn.value, "root_mask": rm.value,
                "capacity": cap.value}

    def raise_for_status(self, bits: int):
        if bits & _lib.S_TABLE_FULL:
            raise _lib.B200RouterError("GPU trie is full (capacity %d nodes): create it "
                                       "with a larger cap_log2" % (1 << self.cap_log2))
        if bits & _lib.S_CHUNK_CAP:
            raise _lib.B200RouterError("chunk capacity exceeded: max_chunks was too small")
        if bits & _lib.S_INTERNAL:
            raise _lib.B200RouterError("internal error: inconsistent trie table")

    def check_status(self, count_nodes: bool = False):
        s = self.status(count_nodes)
        self.raise_for_status(s["status"])
        return s

    def status_async(self, out_word: torch.Tensor):
        """Copy the latched status word into a device int32 in stream order (no host sync)."""
        with torch.cuda.device(self.device):
            _lib.check(_lib.lib().b200r_trie_status_async(self._h, _ptr(out_word), _stream()))

    # -------------------------------------------------------------------- a1
    def hash_chunks(self, p: PackedPrompts) -> Tuple[torch.Tensor, torch.Tensor]:
        """HashTrie._chunk_and_hash for a batch: (chunk_base int32[n+1], hashes int64[...])."""
        p = self._dev(p)
        L = _lib.lib()
        chunk_base = torch.empty(p.n + 1, dtype=torch.int32, device=self.device)
        hashes = torch.empty(max(p.max_chunks, 1), dtype=torch.int64, device=self.device)
        need = L.b200r_xxh64_chunks_scratch_bytes(p.n, p.max_chunks)
        sc = self._scratch_for(need)
        flags = _lib.F_ALL_ASCII if p.all_ascii else 0
        with torch.cuda.device(self.device):
            _lib.check(L.b200r_xxh64_chunks(_ptr(p.bytes), _ptr(p.starts), _ptr(p.lens),
                                            _ptr(p.nchars), p.n, self.chunk_chars, flags,
                                            _ptr(chunk_base), _ptr(hashes), p.max_chunks,
                                            _ptr(sc), sc.numel(), _stream()))
        return chunk_base, hashes

    # -------------------------------------------------------------------- a2
    def longest_prefix_match(self, p: PackedPrompts, avail_mask: int, from_root_endpoints: bool = False):
        """Batch of HashTrie.longest_prefix_match: (depth int32[n], matched int64[n]).

        from_root_endpoints=True gives the gateway picker's variant
        (prefix_aware_picker.go:191-213): a request that matches no chunk gets
        root.endpoints & available instead of `available`; together with prompts packed with
        chunk_by="bytes" this reproduces that picker's trie."""
        chunk_base, hashes = self.hash_chunks(p)
        depth, matched = self.lookup_hashes(hashes, chunk_base, p.n, avail_mask)
        if from_root_endpoints:
            root = self.status()["root_mask"] & avail_mask & M64
            root = root - (1 << 64) if root >= (1 << 63) else root
            matched = torch.where(depth == 0, torch.full_like(matched, root), matched)
        return depth, matched

    def lookup_hashes(self, hashes, chunk_base, n: int, avail_mask: int):
        depth = torch.empty(max(n, 1), dtype=torch.int32, device=self.device)
        matched = torch.empty(max(n, 1), dtype=torch.int64, device=self.device)
        with torch.cuda.device(self.device):
            _lib.check(_lib.lib().b200r_trie_lookup_batch(self._h, _ptr(hashes), _ptr(chunk_base), n,
                                                          avail_mask & M64, _ptr(depth),
                                                          _ptr(matched), _stream()))
        return depth[:n], matched[:n]

    # -------------------------------------------------------------------- a3
    def insert(self, p: PackedPrompts, endpoint_idx: torch.Tensor):
        """Batch of HashTrie.insert (endpoint given as bit index per request)."""
        chunk_base, hashes = self.hash_chunks(p)
        self.insert_hashes(hashes, chunk_base, p.n, endpoint_idx)

    def insert_hashes(self, hashes, chunk_base, n: int, endpoint_idx: torch.Tensor):
        ep = endpoint_idx.to(device=self.device, dtype=torch.int32).contiguous()
        with torch.cuda.device(self.device):
            _lib.check(_lib.lib().b200r_trie_insert_batch(self._h, _ptr(hashes), _ptr(chunk_base), n,
                                                          _ptr(ep), _stream()))

    # -------------------------------------------------------------------- a5
    def route(self, p: PackedPrompts, avail_mask: int, seed: int, seq0: int, out=None):
        """Lookup + tie-break + insert for a batch, equal to routing the requests
        sequentially in index order (b200r_prefix_route_batch).  Returns device tensors
        (endpoint int32[n], depth int32[n], matched int64[n])."""
        p = self._dev(p)
        L = _lib.lib()
        n = p.n
        if out is None:
            out = (torch.empty(max(n, 1), dtype=torch.int32, device=self.device),
                   torch.empty(max(n, 1), dtype=torch.int32, device=self.device),
                   torch.empty(max(n, 1), dtype=torch.int64, device=self.device))
        ep, depth, matched = out
        flags = _lib.F_ALL_ASCII if p.all_ascii else 0
        sc = self._scratch_for(L.b200r_prefix_route_scratch_bytes(n, p.max_chunks, flags))
        with torch.cuda.device(self.device):
            _lib.check(L.b200r_prefix_route_batch(
                self._h, _ptr(p.bytes), _ptr(p.starts), _ptr(p.lens), _ptr(p.nchars), n,
                p.max_chunks, flags, avail_mask & M64, seed & M64, seq0 & M64, _ptr(ep), _ptr(depth),
                _ptr(matched), _ptr(sc), sc.numel(), _stream()))
        return ep[:n], depth[:n], matched[:n]

    # ------------------------------------------------- a5 from chunk hashes
    def route_hashes(self, hashes: torch.Tensor, chunk_base: torch.Tensor, n: int, max_chunks: int,
                     avail_mask: int, seed: int, seq0: int, out=None):
        """The route stage from chunk hashes in CSR form (as hash_chunks returns them):
        b200r_prefix_route_hashes.  Same results as route() on the prompts behind them."""
        L = _lib.lib()
        if out is None:
            out = (torch.empty(max(n, 1), dtype=torch.int32, device=self.device),
                   torch.empty(max(n, 1), dtype=torch.int32, device=self.device),
                   torch.empty(max(n, 1), dtype=torch.int64, device=self.device))
        ep, depth, matched = out
        sc = self._scratch_for(L.b200r_prefix_route_scratch_bytes(n, max_chunks, _lib.F_ALL_ASCII))
        with torch.cuda.device(self.device):
            _lib.check(L.b200r_prefix_route_hashes(
                self._h, _ptr(hashes), _ptr(chunk_base), n, max_chunks, avail_mask & M64,
                seed & M64, seq0 & M64, _ptr(ep), _ptr(depth), _ptr(matched), _ptr(sc), sc.numel(),
                _stream()))
        return ep[:n], depth[:n], matched[:n]

    # ------------------------------------------------- a5 for one request (latency path)
    def route_one(self, prompt: bytes, avail_mask: int, seed: int, seq: int):
        """b200r_prefix_route_one: one launch and one synchronisation for ONE request given as
        host bytes.  Returns (endpoint, depth, matched mask, latched status bits), or None when
        the latency path does not take the prompt (not ASCII, too long): use route()."""
        ep, depth, matched, st = C.c_int32(), C.c_int32(), C.c_uint64(), C.c_uint32()
        with torch.cuda.device(self.device):
            rc = _lib.lib().b200r_prefix_route_one(
                self._h, prompt, len(prompt), avail_mask & M64, seed & M64, seq & M64,
                C.byref(ep), C.byref(depth), C.byref(matched), C.byref(st), _stream())
        if rc == 1:
            return None
        _lib.check(rc)
        return ep.value, depth.value, matched.value, st.value

    # --------------------------------------------------------------- dump etc
    def dump_raw(self):
        """(entries uint64[count,4], root_mask); an entry is {log index | records left in its run
        << 32, parent's log index (0xFFFFFFFF = root) | 1 << 63 on the first node of a run, chunk
        hash, mask}."""
        s = self.status()
        cap = max(int(s["num_nodes"]), 1)
        out = torch.empty(cap * 4, dtype=torch.int64, device=self.device)
        cnt, rm = C.c_int64(), C.c_uint64()
        with torch.cuda.device(self.device):
            _lib.check(_lib.lib().b200r_trie_dump(self._h, _ptr(out), cap, C.byref(cnt), C.byref(rm),
                                                  _stream()))
        if cnt.value > cap:
            raise _lib.B200RouterError("dump overflow: %d > %d" % (cnt.value, cap))
        arr = out.cpu().numpy().view(np.uint64).reshape(-1, 4)[:cnt.value]
        return arr, rm.value

    def restore_entries(self, entries: torch.Tensor, root_mask: int = 0):
        """Put dumped entries back (b200r_trie_restore); the table must be empty."""
        ent = entries.to(self.device).contiguous()
        with torch.cuda.device(self.device):
            _lib.check(_lib.lib().b200r_trie_restore(self._h, _ptr(ent), ent.numel() // 4,
                                                     root_mask & M64, _stream()))

    def retire_endpoints(self, bits: int):
        """Clear the given endpoint bits from every node and from the root (the bit of an
        endpoint that is gone for good can then be handed to a new url)."""
        with torch.cuda.device(self.device):
            _lib.check(_lib.lib().b200r_trie_retire_endpoints(self._h, bits & M64, _stream()))

    # ------------------------------------------------------- checkpoint (f4)
    def save(self, path, extra: Optional[Dict] = None):
        """Write the table to disk (the reference has no checkpoint; its trie dies with the
        process).  Format: npz with entries uint64[count,4] (see dump_raw), root mask, chunk
        size, format version, plus whatever the caller adds in `extra` (the router stores its
        url <-> bit registry there, without which the masks mean nothing to a new process).
        `path` is used as given (no suffix is appended)."""
        arr, root = self.dump_raw()
        fields = dict(entries=arr, root_mask=np.uint64(root), chunk_chars=np.int32(self.chunk_chars),
                      version=np.int32(2))
        for k, v in (extra or {}).items():
            fields["x_" + k] = np.asarray(v)
        with open(path, "wb") as f:
            np.savez(f, **fields)

    @classmethod
    def load(cls, path, cap_log2: Optional[int] = None, device=None, with_extra: bool = False):
        with open(path, "rb") as f:
            z = dict(np.load(f, allow_pickle=False))
        if int(z["version"]) != 2:
            raise _lib.B200RouterError("unknown trie checkpoint version %d" % int(z["version"]))
        arr = z["entries"]
        need = int((arr[:, 0] & np.uint64(0xFFFFFFFF)).max()) + 1 if len(arr) else 1
        if cap_log2 is None:
            cap_log2 = max(12, int(np.ceil(np.log2(need * 1.5))))
        if (1 << cap_log2) < need:
            raise _lib.B200RouterError("checkpoint needs a capacity of at least %d nodes" % need)
        t = cls(cap_log2=cap_log2, chunk_chars=int(z["chunk_chars"]), device=device)
        t.restore_entries(torch.from_numpy(arr.view(np.int64).copy()), int(z["root_mask"]))
        t.check_status()
        if with_extra:
            return t, {k[2:]: v for k, v in z.items() if k.startswith("x_")}
        return t

    def dump(self) -> Dict[tuple, int]:
        """Decode the table back into the reference's shape:
        {tuple(chunk hashes from the root): mask}; () is the root."""
        arr, root = self.dump_raw()
        return decode_dump(arr, root)


ROOT_IDX = 0xFFFFFFFF


def decode_dump(arr: np.ndarray, root_mask: int) -> Dict[tuple, int]:
    """Follow the parent links of a dump up to the root: {hash path: mask}."""
    nodes = {int(e0) & 0xFFFFFFFF: (int(par) & 0xFFFFFFFF, int(h), int(m)) for e0, par, h, m in arr}
    paths: Dict[int, tuple] = {ROOT_IDX: ()}

    def path_of(idx):
        chain = []
        while idx not in paths:
            if idx not in nodes:
                raise ValueError("dump is not prefix-closed: missing parent of a node")
            par, h, _ = nodes[idx]
            chain.append((idx, h))
            idx = par
        base = paths[idx]
        for k, h in reversed(chain):
            base = base + (h,)
            paths[k] = base
        return base

    out = {(): int(root_mask)}
    for idx, (_, _, m) in nodes.items():
        p = path_of(idx)
        if p in out:
            raise ValueError("two nodes with the same path in a dump")
        out[p] = m
    return out


class RoutePipeline:
    """Overlaps consecutive batches on one GPU.

    A routed batch is three steps that meet only in its scratch buffer:
      prep    chunk counts, CSR, tile map, ragged last chunks (latency-bound, a few MB)
      hash    XXH64 of every full chunk: streams the prompt bytes (HBM-bound)
      route   claim / select / commit: walks and extends the table (latency-bound)
    Only `route` touches the table.  The pipeline runs prep of batch b+2 and hash of batch b+1
    on their own streams while route of batch b runs on the caller's stream; route stages stay
    in submission order, so results are still those of sequential routing.  One scratch buffer
    per batch in flight (depth 3).

    Host (pinned) inputs are copied to the device on the prep stream, so the H2D copy of a
    later batch also overlaps the routing of the current one.  For inputs that are already on
    the device pass inputs_ready=True when they are complete (e.g. produced and synchronised
    earlier); otherwise the prep stage waits for the caller's stream, which is safe but
    serialises the stages."""

    def __init__(self, trie: GpuHashTrie, depth: int = 3):
        self.trie = trie
        self.depth = depth
        self.device = trie.device
        self._p = C.c_void_p()
        with torch.cuda.device(self.device):
            _lib.check(_lib.lib().b200r_pipe_create(trie._h, depth, C.byref(self._p)))
        self._scratch = [None] * depth
        self._keep = [None] * depth
        self._outs = [None] * (depth + 1)
        self._i = 0

    def close(self):
        if getattr(self, "_p", None) and self._p.value:
            _lib.lib().b200r_pipe_destroy(self._p)
            self._p = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _scratch_for(self, k, nbytes):
        t = self._scratch[k]
        if t is None or t.numel() < nbytes:
            t = self._scratch[k] = torch.empty(int(nbytes * 1.25) + 4096, dtype=torch.uint8,
                                               device=self.device)
        return t

    def reserve(self, n: int, max_chunks: int, all_ascii: bool = True):
        flags = _lib.F_ALL_ASCII if all_ascii else 0
        need = _lib.lib().b200r_prefix_route_scratch_bytes(n, max_chunks, flags)
        for k in range(self.depth):
            self._scratch_for(k, need)
        for j in range(self.depth + 1):
            self._out_for(j, n)

    def _out_for(self, j, n):
        o = self._outs[j]
        if o is None or o[0].numel() < n:
            o = self._outs[j] = (torch.empty(max(n, 1), dtype=torch.int32, device=self.device),
                                 torch.empty(max(n, 1), dtype=torch.int32, device=self.device),
                                 torch.empty(max(n, 1), dtype=torch.int64, device=self.device))
        return o

    def _copy_on_prep_stream(self, p: PackedPrompts) -> PackedPrompts:
        L = _lib.lib()

        def mv(t):
            if t is None:
                return None
            d = torch.empty(t.shape, dtype=t.dtype, device=self.device)
            _lib.check(L.b200r_pipe_copy_h2d(self._p, d.data_ptr(), t.data_ptr(),
                                             t.numel() * t.element_size()))
            return d
        return PackedPrompts(mv(p.bytes), mv(p.starts), mv(p.lens), mv(p.nchars), p.n, p.max_chunks,
                             p.all_ascii, p.payload_bytes)

    def route(self, p: PackedPrompts, avail_mask: int, seed: int, seq0: int, out=None,
              inputs_ready: bool = False):
        """Returns device tensors (endpoint, depth, matched), valid in stream order on
        the caller's current stream.  Unless `out` is given they live in a ring of
        depth+1 buffers: consume them before depth+1 further batches are submitted.
        One native call per batch (b200r_pipe_route)."""
        L = _lib.lib()
        n = p.n
        if n == 0:
            e = torch.empty(0, dtype=torch.int32, device=self.device)
            return e, e.clone(), torch.empty(0, dtype=torch.int64, device=self.device)
        k = self._i % self.depth
        o = out if out is not None else self._out_for(self._i % (self.depth + 1), n)
        self._i += 1
        flags = _lib.F_ALL_ASCII if p.all_ascii else 0
        sc = self._scratch[k]
        need = L.b200r_prefix_route_scratch_bytes(n, p.max_chunks, flags)
        if sc is None or sc.numel() < need:
            sc = self._scratch_for(k, need)
        on_prep = inputs_ready
        cur = torch.cuda.current_stream(self.device)
        if p.bytes.device.type == "cpu":
            # copy on the prep stream, after the call that used this rotation slot has finished;
            # its device copy of the inputs is released only now, so the allocator can hand the
            # memory to this copy without overtaking a kernel that still reads it
            _lib.check(L.b200r_pipe_begin(self._p))
            self._keep[k] = None
            with torch.cuda.device(self.device):
                p = self._copy_on_prep_stream(p)
            self._keep[k] = p
            on_prep = True
        ep, depth, matched = o
        rc = L.b200r_pipe_route(self._p, p.bytes.data_ptr(), p.starts.data_ptr(), p.lens.data_ptr(),
                                None if p.nchars is None else p.nchars.data_ptr(), n, p.max_chunks,
                                flags, avail_mask & M64, seed & M64, seq0 & M64, ep.data_ptr(),
                                depth.data_ptr(), matched.data_ptr(), sc.data_ptr(), sc.numel(),
                                1 if on_prep else 0, cur.cuda_stream)
        if rc:
            _lib.check(rc)
        if ep.shape[0] == n:                      # a tensor slice costs microseconds
            return ep, depth, matched
        return ep[:n], depth[:n], matched[:n]

    def check_status(self, count_nodes: bool = False):
        return self.trie.check_status(count_nodes)
