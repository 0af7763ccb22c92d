"""Prefix-aware routing of one global batch sharded over the GPUs of a box.

One process per GPU (torch.distributed, NCCL over NVLink / NVSwitch).  The global batch is
the rank-major concatenation of the ranks' shards and must be routed exactly as if the
requests arrived one after another in that order.  Two designs:

ShardedPrefixRouter -- table REPLICATED per GPU (north_star's wording).  Every rank hashes
  its own shard (the HBM-bound part, 1/world of the work), the ranks all-gather what defines
  the batch's new trie nodes -- the chunk hashes, 8 bytes per 128-byte chunk -- and every
  rank routes the whole global batch against its replica with the ordinary route stage
  (b200r_prefix_route_hashes), so all replicas receive all inserts and stay equal in
  content.  Per-GPU routing work grows with the world size: it cannot weak-scale.

PartitionedPrefixRouter -- table PARTITIONED by first-chunk hash, chunk hashes shipped to
  the owning rank over NVLink peer memory by the library's own kernels; per-GPU work stays
  constant (the default of bench.py --gpus N).

`ops` does the device work.  The product uses CudaShardOps (C ABI of libb200router.so, no
CPU fallback); the gloo test in tests/ injects an emulation so that the orchestration is
exercised with world_size 2 on CPUs.
"""
import ctypes as C

import torch
import torch.distributed as dist

from . import _lib
from .packing import PackedPrompts

M64 = (1 << 64) - 1


def _ptr(t):
    return None if t is None else C.c_void_p(t.data_ptr())


def _stream():
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


class CudaShardOps:
    """The two device steps of the replicated design, through the C ABI."""

    def __init__(self, cap_log2: int, device, chunk_chars: int = 128):
        from .trie import GpuHashTrie
        self.trie = GpuHashTrie(cap_log2=cap_log2, chunk_chars=chunk_chars, device=device)
        self.device = self.trie.device

    def hash_chunks(self, p: PackedPrompts):
        """(chunk_base int32[n+1], hashes int64[>= chunk_base[n]]) on the device."""
        return self.trie.hash_chunks(p)

    def route_hashes(self, hashes, chunk_base, n, max_chunks, avail_mask, seed, seq0):
        return self.trie.route_hashes(hashes, chunk_base, n, max_chunks, avail_mask, seed, seq0)

    def check_status(self, count_nodes=False):
        return self.trie.check_status(count_nodes)

    def dump(self):
        return self.trie.dump()


class TorchComm:
    """all-gather over a torch.distributed process group (NCCL on GPUs, gloo in CPU tests)."""

    def __init__(self, group=None):
        self.group = group
        self.world = dist.get_world_size(group)
        self.rank = dist.get_rank(group)

    def all_gather(self, out_flat: torch.Tensor, inp_flat: torch.Tensor):
        dist.all_gather_into_tensor(out_flat, inp_flat, group=self.group)


class ShardedPrefixRouter:
    """route(shard) on every rank == sequential routing of the concatenated batch."""

    def __init__(self, cap_log2: int = 22, device=None, max_requests: int = 0, max_chunks: int = 0,
                 group=None, ops=None, comm=None):
        self.comm = comm if comm is not None else TorchComm(group)
        self.world = self.comm.world
        self.rank = self.comm.rank
        self.ops = ops if ops is not None else CudaShardOps(cap_log2, device)
        self.device = self.ops.device
        self.last_exchange_bytes = 0

    def route(self, p: PackedPrompts, avail_mask: int, seed: int, seq0: int, out=None):
        """seq0 = global sequence number of the batch's first request (same on all ranks)."""
        if p.bytes.device != self.device:
            p = p.to(self.device)
        W, rank, n = self.world, self.rank, p.n
        chunk_base, hashes = self.ops.hash_chunks(p)
        # gather #1: shard sizes {requests, chunks}; the one host read of the batch
        meta = torch.stack([torch.tensor(n, dtype=torch.int32, device=self.device),
                            chunk_base[n].to(torch.int32)])
        meta_all = torch.empty(W * 2, dtype=torch.int32, device=self.device)
        self.comm.all_gather(meta_all, meta)
        mh = meta_all.cpu().view(W, 2)
        ns, cs = mh[:, 0].tolist(), mh[:, 1].tolist()
        nmax, cmax = max(max(ns), 1), max(max(cs), 1)
        # gather #2 / #3: per-request chunk counts and the chunk hashes, padded to the largest shard
        counts = torch.zeros(nmax, dtype=torch.int32, device=self.device)
        counts[:n] = chunk_base[1:n + 1] - chunk_base[:n]
        hpad = torch.zeros(cmax, dtype=torch.int64, device=self.device)
        hpad[:cs[rank]] = hashes[:cs[rank]]
        counts_all = torch.empty(W * nmax, dtype=torch.int32, device=self.device)
        hashes_all = torch.empty(W * cmax, dtype=torch.int64, device=self.device)
        self.comm.all_gather(counts_all, counts)
        self.comm.all_gather(hashes_all, hpad)
        # the global batch in rank-major order, as CSR
        g_counts = torch.cat([counts_all[q * nmax:q * nmax + ns[q]] for q in range(W)])
        g_hashes = torch.cat([hashes_all[q * cmax:q * cmax + cs[q]] for q in range(W)] +
                             [torch.zeros(1, dtype=torch.int64, device=self.device)])
        N, Ctot = sum(ns), sum(cs)
        g_base = torch.zeros(N + 1, dtype=torch.int32, device=self.device)
        g_base[1:] = torch.cumsum(g_counts, 0)
        ep, depth, matched = self.ops.route_hashes(g_hashes, g_base, N, Ctot, avail_mask, seed, seq0)
        off = sum(ns[:rank])
        res = (ep[off:off + n], depth[off:off + n], matched[off:off + n])
        if out is not None:
            for o, r in zip(out, res):
                o[:n].copy_(r)
            res = tuple(o[:n] for o in out)
        self.last_exchange_bytes = (W - 1) * (cmax * 8 + nmax * 4 + 8)
        return res

    def check_status(self, count_nodes=False):
        return self.ops.check_status(count_nodes)

    def dump(self):
        return self.ops.dump()


PART_SETS = 3        # buffer sets of the partitioned exchange (csrc/internal.cuh)


class PartitionedPrefixRouter:
    """Multi-GPU routing with a PARTITIONED table -- the variant that weak-scales.

    A request's whole path hangs under its first chunk, so requests whose first chunk
    hashes differ never meet in the trie.  Rank owner(first chunk hash) % world holds
    that subtree; nothing is replicated.  Per batch every rank hashes its own shard,
    pushes the chunk hashes of each request (8 bytes per 128-byte chunk) to the owning
    rank over NVLink peer memory, routes the requests it owns -- in global order --
    against its partition with the ordinary claim / select / commit kernels, and pushes
    the results back.  No NCCL call and no host read per batch; the tie-break sequence
    number lives on the device (starts at 0).  Same results as sequential routing of the
    rank-major global batch; the union of the partitions is the reference trie.
    (Deviation from north_star's "replicated table": see DESIGN.md 4.)"""

    def __init__(self, cap_log2: int = 22, device=None, max_requests: int = 0, max_chunks: int = 0,
                 group=None, _sim=None):
        from .trie import GpuHashTrie
        L = _lib.lib()
        if _sim is None:
            self.world, self.rank = dist.get_world_size(group), dist.get_rank(group)
        else:
            self.world, self.rank = _sim["world"], _sim["rank"]
        self.trie = GpuHashTrie(cap_log2=cap_log2, device=device)
        self.device = self.trie.device
        if _sim is None and self.world > 1:
            caps = torch.tensor([int(max_requests), int(max_chunks)], dtype=torch.int64,
                                device=self.device)
            dist.all_reduce(caps, op=dist.ReduceOp.MAX, group=group)
            max_requests, max_chunks = (int(x) for x in caps.tolist())
        self.cap, self.rcap = self.capacities(max_requests, max_chunks)
        lay = (C.c_size_t * 2)()
        _lib.check(L.b200r_part_layout(self.world, self.cap, self.rcap, lay))
        self.symm_bytes, work_bytes = int(lay[0]), int(lay[1])
        if _sim is None:
            import torch.distributed._symmetric_memory as symm_mem
            self.symm = symm_mem.empty(self.symm_bytes, dtype=torch.uint8, device=self.device)
            self.symm.zero_()
            torch.cuda.synchronize(self.device)
            g = group if group is not None else dist.group.WORLD
            self._hdl = symm_mem.rendezvous(self.symm, g.group_name)
            bases = list(self._hdl.buffer_ptrs)
            self._hdl.barrier()
        else:
            self.symm = _sim["buffers"][self.rank]
            bases = [b.data_ptr() for b in _sim["buffers"]]
        self.peer_bases = torch.tensor(bases, dtype=torch.int64, device=self.device)
        self.work = torch.zeros(work_bytes, dtype=torch.uint8, device=self.device)
        self.epoch = 0
        self._shape = None
        # route() pipelines consecutive batches natively (b200r_part_step); the phase_* methods
        # run the three phases one by one on the current stream (lock-step tests)
        self.pipelined = True
        self._keep = [None] * PART_SETS            # device copies of host inputs, by buffer set
        self._sc = [None] * PART_SETS              # hash-stage scratch, by buffer set
        self._retired = []                         # outgrown scratch buffers (may still be in use)
        self._sc_lim = [None] * PART_SETS          # (n, max_chunks, flags) each scratch buffer is good for
        self._const_ptrs = None

    @staticmethod
    def capacities(max_requests: int, max_chunks: int):
        # even capacities: the exchange ships 16-byte vectors into per-source regions
        return (max(int(max_chunks), 64) + 1) // 2 * 2, (max(int(max_requests), 4) + 1) // 2 * 2

    @staticmethod
    def symm_bytes_for(world: int, max_requests: int, max_chunks: int) -> int:
        cap, rcap = PartitionedPrefixRouter.capacities(max_requests, max_chunks)
        lay = (C.c_size_t * 2)()
        _lib.check(_lib.lib().b200r_part_layout(world, cap, rcap, lay))
        return int(lay[0])

    def phase_send(self, p: PackedPrompts):
        if p.bytes.device != self.device:
            p = p.to(self.device)
        self.epoch += 1
        flags = _lib.F_ALL_ASCII if p.all_ascii else 0
        need = _lib.lib().b200r_prefix_route_scratch_bytes(p.n, p.max_chunks, flags)
        sc = self.trie._scratch_for(need)
        self._shape = (p.n,)
        with torch.cuda.device(self.device):
            _lib.check(_lib.lib().b200r_part_send(
                self.trie._h, _ptr(p.bytes), _ptr(p.starts), _ptr(p.lens), _ptr(p.nchars), p.n,
                p.max_chunks, flags, _ptr(self.peer_bases), self.world, self.rank, self.cap,
                self.rcap, self.epoch, _ptr(self.work), _ptr(sc), sc.numel(), _stream()))

    def phase_route(self, avail_mask: int, seed: int):
        with torch.cuda.device(self.device):
            _lib.check(_lib.lib().b200r_part_route(
                self.trie._h, avail_mask & M64, seed & M64, _ptr(self.peer_bases), _ptr(self.symm),
                self.world, self.rank, self.cap, self.rcap, self.epoch, _ptr(self.work), _stream()))

    def phase_finish(self, out=None):
        n = self._shape[0]
        if out is None:
            out = (torch.empty(max(n, 1), dtype=torch.int32, device=self.device),
                   torch.empty(max(n, 1), dtype=torch.int32, device=self.device),
                   torch.empty(max(n, 1), dtype=torch.int64, device=self.device))
        ep, depth, matched = out
        with torch.cuda.device(self.device):
            _lib.check(_lib.lib().b200r_part_finish(
                self.trie._h, n, _ptr(self.symm), self.world, self.rank, self.cap, self.rcap,
                self.epoch, _ptr(self.work), _ptr(ep), _ptr(depth), _ptr(matched), _stream()))
        return ep[:n], depth[:n], matched[:n]

    def route(self, p: PackedPrompts, avail_mask: int, seed: int, seq0: int = 0, out=None,
              inputs_ready: bool = False):
        """One native call per batch (b200r_part_step): prep, send (hashing + shipment), route and
        return run on four streams of the library and the caller's stream waits for the results.
        Send of batch b+1 overlaps route + return of batch b, the route chain of batch b+1 starts
        right behind the commit of batch b.  The send may not start before this rank has finished
        batch b-1 (then every peer has consumed the buffer set it is about to overwrite) and,
        unless inputs_ready, before the caller's stream has produced the inputs.  Host (pinned)
        inputs are copied on the prep stream."""
        if self.epoch == 0 and seq0 != 0:
            raise _lib.B200RouterError("PartitionedPrefixRouter keeps the sequence number on the "
                                       "device; it starts at 0")
        L = _lib.lib()
        self.epoch += 1
        n = p.n
        # the wrapper is on the critical path of every batch (the step is bound by host time):
        # constant pointers are cached, the device is switched only when it is not current
        dev_idx = self.device.index if self.device.index is not None else torch.cuda.current_device()
        switch = torch.cuda.current_device() != dev_idx
        if switch:
            prev_dev = torch.cuda.current_device()
            torch.cuda.set_device(dev_idx)
        try:
            if p.bytes.device != self.device:
                _lib.check(L.b200r_part_begin(self.trie._h, self.epoch))
                k = self.epoch % PART_SETS
                self._keep[k] = None
                self._keep[k] = p = self._copy_on_prep_stream(p)
                inputs_ready = True
            flags = _lib.F_ALL_ASCII if p.all_ascii else 0
            sc = self._step_scratch_for(n, p.max_chunks, flags)
            self._shape = (n,)
            if out is None:
                out = (torch.empty(max(n, 1), dtype=torch.int32, device=self.device),
                       torch.empty(max(n, 1), dtype=torch.int32, device=self.device),
                       torch.empty(max(n, 1), dtype=torch.int64, device=self.device))
            ep, depth, matched = out
            c = self._const_ptrs
            if c is None:
                c = self._const_ptrs = (self.trie._h, self.peer_bases.data_ptr(), self.symm.data_ptr(),
                                        self.work.data_ptr())
            rc = L.b200r_part_step(
                c[0], p.bytes.data_ptr(), p.starts.data_ptr(), p.lens.data_ptr(),
                None if p.nchars is None else p.nchars.data_ptr(), n, p.max_chunks, flags,
                avail_mask & M64, seed & M64, c[1], c[2],
                self.world, self.rank, self.cap, self.rcap, self.epoch, c[3],
                sc.data_ptr(), sc.numel(), ep.data_ptr(), depth.data_ptr(), matched.data_ptr(),
                1 if inputs_ready else 0, torch.cuda.current_stream().cuda_stream)
            if rc:
                _lib.check(rc)
        finally:
            if switch:
                torch.cuda.set_device(prev_dev)
        if ep.shape[0] == n:                      # a tensor slice costs microseconds
            return ep, depth, matched
        return ep[:n], depth[:n], matched[:n]

    def _step_scratch_for(self, n: int, max_chunks: int, flags: int) -> torch.Tensor:
        # the size is monotone in n and max_chunks: ask the library only when either grew
        k = self.epoch % PART_SETS
        lim = self._sc_lim[k]
        if lim is not None and n <= lim[0] and max_chunks <= lim[1] and flags == lim[2]:
            return self._sc[k]
        need = _lib.lib().b200r_prefix_route_scratch_bytes(n, max_chunks, flags)
        sc = self._step_scratch(need)
        self._sc_lim[k] = (n, max_chunks, flags)
        return sc

    def _step_scratch(self, nbytes: int) -> torch.Tensor:
        # one buffer per set in turn: the prep of batch b+1 runs next to the hashing of batch b
        k = self.epoch % PART_SETS
        sc = self._sc[k]
        if sc is None or sc.numel() < nbytes:
            if sc is not None:
                self._retired.append(sc)
            sc = self._sc[k] = torch.empty(int(nbytes * 1.25) + 4096, dtype=torch.uint8,
                                           device=self.device)
        return sc

    def _copy_on_prep_stream(self, p: PackedPrompts) -> PackedPrompts:
        L = _lib.lib()

        def mv(t):
            if t is None:
                return None
            d = torch.empty(t.shape, dtype=t.dtype, device=self.device)
            _lib.check(L.b200r_part_copy_h2d(self.trie._h, d.data_ptr(), t.data_ptr(),
                                             t.numel() * t.element_size()))
            return d
        return PackedPrompts(mv(p.bytes), mv(p.starts), mv(p.lens), mv(p.nchars), p.n, p.max_chunks,
                             p.all_ascii, p.payload_bytes)

    def check_status(self, count_nodes=False):
        s = self.trie.check_status(count_nodes)
        if s["status"] & _lib.S_PEER_TIMEOUT:
            raise _lib.B200RouterError("a peer GPU did not deliver its part of a batch in time")
        return s

    def dump(self):
        """This rank's partition; the union over ranks (root masks OR-ed) is the whole trie."""
        return self.trie.dump()
