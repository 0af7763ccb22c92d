"""Prefix-aware routing of one global batch sharded over the GPUs of a box.

One process per GPU (torch.distributed, NCCL over NVLink / NVSwitch).  Every
rank keeps a full replica of the trie table; the global batch is the rank-major
concatenation of the ranks' shards and must be routed exactly as if the
requests arrived one after another in that order.

Per batch, on every rank:
  classify   hash the own shard, look every chunk up in the (identical) local
             replica; chunks whose node is not usable yet become 32-byte records
  gather #1  shard sizes + record counts (one tiny all-gather, one host read)
  gather #2  the records themselves
  apply      all ranks' records -> find-or-create + global "first toucher"
  select     own shard: deepest usable node, root draw or leader (global order)
  gather #3  leaders / choices / match depths of all shards
  commit     resolve every request's endpoint; OR it into the new nodes of
             every replica; write the own shard's results

`ops` does the arithmetic.  The product uses CudaShardOps (C ABI of
libb200router.so, no CPU fallback); the gloo test in tests/ injects an
emulation so that this orchestration is exercised with world_size 2 on CPUs.
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
    """The four per-batch device steps, through the C ABI (include/b200_router.h)."""

    def __init__(self, cap_log2: int, device, chunk_chars: int = 128):
        from .trie import GpuHashTrie
        self.trie = GpuHashTrie(cap_log2=cap_log2, chunk_chars=chunk_chars, device=device)
        self.device = self.trie.device
        self._shape = None

    def _scratch(self, p: PackedPrompts):
        flags = _lib.F_ALL_ASCII if p.all_ascii else 0
        need = _lib.lib().b200r_prefix_route_scratch_bytes(p.n, p.max_chunks, flags)
        self._shape = (p.n, p.max_chunks, flags)
        return self.trie._scratch_for(need), flags

    def classify(self, p: PackedPrompts, avail_mask: int, recs: torch.Tensor, count: torch.Tensor):
        sc, flags = self._scratch(p)
        with torch.cuda.device(self.device):
            _lib.check(_lib.lib().b200r_shard_classify(
                self.trie._h, _ptr(p.bytes), _ptr(p.starts), _ptr(p.lens), _ptr(p.nchars), p.n,
                p.max_chunks, flags, avail_mask & M64, _ptr(recs), _ptr(count), _ptr(sc), sc.numel(),
                _stream()))

    def apply(self, recs_all, meta_all, world, stride, rank, rec_slot):
        n, max_chunks, flags = self._shape
        with torch.cuda.device(self.device):
            _lib.check(_lib.lib().b200r_shard_apply(
                self.trie._h, _ptr(recs_all), _ptr(meta_all), world, stride, rank, n, max_chunks,
                flags, _ptr(rec_slot), _ptr(self.trie._scratch), _stream()))

    def select(self, avail_mask, seed, seq0, my_off, leader, choice, depth, matched):
        n, max_chunks, flags = self._shape
        with torch.cuda.device(self.device):
            _lib.check(_lib.lib().b200r_shard_select(
                self.trie._h, n, max_chunks, flags, avail_mask & M64, seed & M64, seq0 & M64, my_off,
                _ptr(leader), _ptr(choice), _ptr(depth), _ptr(matched), _ptr(self.trie._scratch),
                _stream()))

    def commit(self, recs_all, meta_all, world, stride, rec_slot, lcd_all, rstride, rank,
               endpoint_all, endpoint, depth, matched):
        n = self._shape[0]
        with torch.cuda.device(self.device):
            _lib.check(_lib.lib().b200r_shard_commit(
                self.trie._h, _ptr(recs_all), _ptr(meta_all), world, stride, _ptr(rec_slot),
                _ptr(lcd_all), rstride, rank, n, _ptr(endpoint_all), _ptr(endpoint), _ptr(depth),
                _ptr(matched), _stream()))

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
        self._recs = None          # int64 [cap, 4]: this rank's records
        self._count = torch.zeros(2, dtype=torch.int32, device=self.device)   # {records, n}
        self._bufs = {}
        if max_chunks:
            self._ensure_recs(max_chunks)
        self.last_exchange_bytes = 0

    def _ensure_recs(self, cap):
        if self._recs is None or self._recs.shape[0] < cap:
            old = self._recs
            self._recs = torch.empty((int(cap * 1.1) + 64, 4), dtype=torch.int64, device=self.device)
            if old is not None:
                self._recs[:old.shape[0]] = old
        return self._recs

    def _buf(self, name, shape, dtype):
        t = self._bufs.get(name)
        numel = 1
        for s in shape:
            numel *= s
        if t is None or t.numel() < numel or t.dtype != dtype:
            t = torch.empty(int(numel * 1.1) + 64, dtype=dtype, device=self.device)
            self._bufs[name] = t
        return t[:numel].view(*shape)

    def route(self, p: PackedPrompts, avail_mask: int, seed: int, seq0: int, out=None):
        """seq0 = global sequence number of the batch's first request (same on all ranks)."""
        if p.bytes.device != self.device:
            p = p.to(self.device)
        W, rank, n = self.world, self.rank, p.n
        recs = self._ensure_recs(max(p.max_chunks, 1))
        self.ops.classify(p, avail_mask, recs, self._count)

        # gather #1: record counts and shard sizes ({count, n} per rank, written by classify)
        meta_all = self._buf("meta", (W, 2), torch.int32)
        self.comm.all_gather(meta_all.view(-1), self._count)
        meta_host = meta_all.cpu()                       # the one host read of the batch
        cnts = meta_host[:, 0].tolist()
        ns = meta_host[:, 1].tolist()
        stride = max(cnts)
        rstride = max(max(ns), 1)
        my_off = sum(ns[:rank])

        # gather #2: records (rank q at q * stride)
        recs_all = self._buf("recs_all", (W, max(stride, 1), 4), torch.int64)
        rec_slot = self._buf("rec_slot", (W * max(stride, 1),), torch.int32)
        if stride > 0:
            recs = self._ensure_recs(stride)
            self.comm.all_gather(recs_all.view(-1), recs[:stride].reshape(-1))
            self.ops.apply(recs_all, meta_all, W, stride, rank, rec_slot)

        # select on the own shard; gather #3: leaders, choices, match depths
        lcd = self._buf("lcd", (3, rstride), torch.int32)
        if out is None:
            out = (torch.empty(max(n, 1), dtype=torch.int32, device=self.device),
                   torch.empty(max(n, 1), dtype=torch.int32, device=self.device),
                   torch.empty(max(n, 1), dtype=torch.int64, device=self.device))
        ep, depth, matched = out
        self.ops.select(avail_mask, seed, seq0, my_off, lcd[0], lcd[1], lcd[2], matched)
        lcd_all = self._buf("lcd_all", (W, 3, rstride), torch.int32)
        self.comm.all_gather(lcd_all.view(-1), lcd.reshape(-1))
        endpoint_all = self._buf("ep_all", (W * rstride,), torch.int32)
        self.ops.commit(recs_all, meta_all, W, stride, rec_slot, lcd_all, rstride, rank,
                        endpoint_all, ep, depth, matched)
        self.last_exchange_bytes = (W - 1) * (stride * 32 + rstride * 12 + 8)
        return ep[:n], depth[:n], matched[:n]

    def check_status(self, count_nodes=False):
        return self.ops.check_status(count_nodes)

    def dump(self):
        return self.ops.dump()


class P2PShardedRouter:
    """Same global-batch semantics as ShardedPrefixRouter, but the exchange is done by
    the library's own kernels over NVLink peer memory: every rank pushes its records
    (and later its leaders / choices / match depths) straight into buffers of all its
    peers and publishes arrival flags; consumers spin on the device.  No NCCL call and
    no host read per batch -- the host only enqueues kernels.

    The peer-addressable buffer comes from torch.distributed._symmetric_memory.  The
    global tie-break sequence number lives on the device (it starts at 0 and advances
    by the global batch size), so `seq0` of route() is only checked on the first batch.
    """

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
        if _sim is None and self.world > 1:           # capacities must agree on all ranks
            caps = torch.tensor([int(max_requests), int(max_chunks)], dtype=torch.int64,
                                device=self.device)
            dist.all_reduce(caps, op=dist.ReduceOp.MAX, group=group)
            max_requests, max_chunks = (int(x) for x in caps.tolist())
        self.cap = max(int(max_chunks), 64)
        self.rcap = (max(int(max_requests), 4) + 3) // 4 * 4
        lay = (C.c_size_t * 3)()
        _lib.check(L.b200r_p2p_layout(self.world, self.cap, self.rcap, lay))
        self.symm_bytes = int(lay[0])
        if _sim is None:
            import torch.distributed._symmetric_memory as symm_mem
            self.symm = symm_mem.empty(self.symm_bytes, dtype=torch.uint8, device=self.device)
            self.symm.zero_()
            torch.cuda.synchronize(self.device)
            g = group if group is not None else dist.group.WORLD
            self._hdl = symm_mem.rendezvous(self.symm, g.group_name)
            bases = list(self._hdl.buffer_ptrs)
            self._hdl.barrier()                       # every peer's flag block is zeroed
        else:                                         # single-device lock-step emulation (tests)
            self.symm = _sim["buffers"][self.rank]
            bases = [b.data_ptr() for b in _sim["buffers"]]
        self.peer_bases = torch.tensor(bases, dtype=torch.int64, device=self.device)
        self.dyn = torch.zeros(1024, dtype=torch.uint8, device=self.device)
        self.recs_local = torch.empty((self.cap, 4), dtype=torch.int64, device=self.device)
        self.rec_count = torch.zeros(2, dtype=torch.int32, device=self.device)
        self.rec_slot = torch.empty(self.world * self.cap, dtype=torch.int32, device=self.device)
        self.lcd_local = torch.zeros(3 * self.rcap, dtype=torch.int32, device=self.device)
        self.endpoint_all = torch.empty(self.world * self.rcap, dtype=torch.int32, device=self.device)
        self.epoch = 0
        self._shape = None
        self._out = None

    @staticmethod
    def symm_bytes_for(world: int, max_requests: int, max_chunks: int) -> int:
        lay = (C.c_size_t * 3)()
        _lib.check(_lib.lib().b200r_p2p_layout(world, max(int(max_chunks), 64),
                                               (max(int(max_requests), 4) + 3) // 4 * 4, lay))
        return int(lay[0])

    # the three phases of a batch (route() runs them back to back)
    def phase_classify(self, p: PackedPrompts, avail_mask: int):
        if p.bytes.device != self.device:
            p = p.to(self.device)
        self.epoch += 1
        flags = _lib.F_ALL_ASCII if p.all_ascii else 0
        need = _lib.lib().b200r_prefix_route_scratch_bytes(p.n, p.max_chunks, flags)
        sc = self.trie._scratch_for(need)
        self._shape = (p.n, p.max_chunks, flags, avail_mask & M64)
        with torch.cuda.device(self.device):
            _lib.check(_lib.lib().b200r_p2p_classify(
                self.trie._h, _ptr(p.bytes), _ptr(p.starts), _ptr(p.lens), _ptr(p.nchars), p.n,
                p.max_chunks, flags, avail_mask & M64, _ptr(self.peer_bases), self.world, self.rank,
                self.cap, self.rcap, self.epoch, _ptr(self.recs_local), _ptr(self.rec_count),
                _ptr(sc), sc.numel(), _stream()))

    def phase_route(self, seed: int, out=None):
        n, max_chunks, flags, avail = self._shape
        if out is None:
            out = (torch.empty(max(n, 1), dtype=torch.int32, device=self.device),
                   torch.empty(max(n, 1), dtype=torch.int32, device=self.device),
                   torch.empty(max(n, 1), dtype=torch.int64, device=self.device))
        self._out = out
        with torch.cuda.device(self.device):
            _lib.check(_lib.lib().b200r_p2p_route(
                self.trie._h, n, max_chunks, flags, avail, seed & M64, _ptr(self.peer_bases),
                _ptr(self.symm), self.world, self.rank, self.cap, self.rcap, self.epoch,
                _ptr(self.dyn), _ptr(self.rec_slot), _ptr(self.lcd_local), _ptr(out[2]),
                _ptr(self.trie._scratch), _stream()))

    def phase_commit(self):
        n = self._shape[0]
        ep, depth, matched = self._out
        with torch.cuda.device(self.device):
            _lib.check(_lib.lib().b200r_p2p_commit(
                self.trie._h, n, _ptr(self.symm), self.world, self.rank, self.cap, self.rcap,
                self.epoch, _ptr(self.dyn), _ptr(self.rec_slot), _ptr(self.endpoint_all), _ptr(ep),
                _ptr(depth), _ptr(matched), _stream()))
        return ep[:n], depth[:n], matched[:n]

    def route(self, p: PackedPrompts, avail_mask: int, seed: int, seq0: int = 0, out=None):
        if self.epoch == 0 and seq0 != 0:
            raise _lib.B200RouterError("P2PShardedRouter keeps the sequence number on the device; "
                                       "it starts at 0")
        self.phase_classify(p, avail_mask)
        self.phase_route(seed, out)
        return self.phase_commit()

    def check_status(self, count_nodes=False):
        s = self.trie.check_status(count_nodes)
        if s["status"] & _lib.S_PEER_TIMEOUT:
            raise _lib.B200RouterError("a peer GPU did not deliver its part of a batch in time")
        return s

    def dump(self):
        return self.trie.dump()


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
        # two-stream pipeline: the send phase of batch b+1 (hashing, never touches the table)
        # runs on its own stream while route + finish of batch b run on the caller's stream
        self.send_stream = torch.cuda.Stream(device=self.device)
        self._sent = torch.cuda.Event()
        self._finished = [None, None]              # finish events by batch parity

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
        """Pipelined: the send phase runs on a second stream.  It may not start before this
        rank has finished batch b-1 (then every peer has consumed the buffer set it is about
        to overwrite) and, unless inputs_ready, before the caller's stream has produced the
        inputs; route and finish stay on the caller's stream, in batch order."""
        if self.epoch == 0 and seq0 != 0:
            raise _lib.B200RouterError("PartitionedPrefixRouter keeps the sequence number on the "
                                       "device; it starts at 0")
        cur = torch.cuda.current_stream(self.device)
        ss = self.send_stream
        prev = self._finished[(self.epoch + 1) & 1]      # batch b-1 (same parity as b+1)
        if prev is not None:
            ss.wait_event(prev)
        if p.bytes.device != self.device:
            inputs_ready = True                          # copied on the send stream below
        if not inputs_ready:
            ss.wait_stream(cur)
        with torch.cuda.stream(ss):
            self.phase_send(p)
            self._sent.record(ss)
        cur.wait_event(self._sent)
        self.phase_route(avail_mask, seed)
        res = self.phase_finish(out)
        done = torch.cuda.Event()
        done.record(cur)
        self._finished[self.epoch & 1] = done
        return res

    def check_status(self, count_nodes=False):
        s = self.trie.check_status(count_nodes)
        if s["status"] & _lib.S_PEER_TIMEOUT:
            raise _lib.B200RouterError("a peer GPU did not deliver its part of a batch in time")
        return s

    def dump(self):
        """This rank's partition; the union over ranks (root masks OR-ed) is the whole trie."""
        return self.trie.dump()
