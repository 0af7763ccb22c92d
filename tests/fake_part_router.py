"""CPU emulation of the PARTITIONED multi-GPU design (variant 3) for the gloo test.

What the product does with kernels over NVLink peer memory (csrc/prefix.cu: k_part_*) is done
here with torch.distributed object collectives and the C oracle, so that the ALGORITHM --
ownership by first-chunk hash, routing each partition's requests in global order with the
tie-break counter taken from the request's global position, results shipped back -- runs in
real separate processes on CPUs.  TEST INFRASTRUCTURE ONLY.
"""
import numpy as np
import torch.distributed as dist

from oracle import coracle

M64 = (1 << 64) - 1


def part_owner(first_chunk_hash, world: int) -> int:
    """csrc/prefix.cu:part_owner -- mix(h0) % world; an empty prompt goes to rank 0."""
    if first_chunk_hash is None:
        return 0
    x = (int(first_chunk_hash) * 0x9E3779B97F4A7C15) & M64
    return (x >> 33) % world


class FakePartitionedRouter:
    def __init__(self, chunk_chars: int = 128):
        self.world, self.rank = dist.get_world_size(), dist.get_rank()
        self.trie = coracle.TrieOracle(chunk_chars)       # this rank's partition
        self.chunk_chars = chunk_chars
        self.seq_base = 0                                 # global sequence number of request 0

    def route(self, prompts, avail_mask: int, seed: int):
        W, me = self.world, self.rank
        # send: per owner, my requests it owns, order preserved, with their local index
        blobs = [p.encode("utf-8") for p in prompts]
        out = [[] for _ in range(W)]
        for i, b in enumerate(blobs):
            hs = coracle.xxh64_chunks(b, self.chunk_chars)
            out[part_owner(hs[0] if len(hs) else None, W)].append((i, b))
        sizes = [None] * W
        dist.all_gather_object(sizes, len(prompts))
        offs = np.concatenate([[0], np.cumsum(sizes)])    # global position of each shard
        inbox = [None] * W
        dist.all_gather_object(inbox, out)                # inbox[q][me] = what rank q sends me
        # route: my requests in global order = sources in rank order, each in its own order
        results = [[] for _ in range(W)]
        for q in range(W):
            for i, b in inbox[q][me]:
                pk = coracle.Packed([b])
                ep, depth, matched = self.trie.route_batch(pk.bytes, pk.starts, pk.lens, avail_mask, seed,
                                                           self.seq_base + int(offs[q]) + i)
                results[q].append((i, int(ep[0]), int(depth[0]), int(matched[0])))
        self.seq_base += int(offs[W])
        # finish: results back to the sources, scattered to request order
        back = [None] * W
        dist.all_gather_object(back, results)
        ep = [None] * len(prompts)
        depth = [None] * len(prompts)
        matched = [None] * len(prompts)
        for o in range(W):
            for i, e, d, m in back[o][me]:
                ep[i], depth[i], matched[i] = e, d, m
        return ep, depth, matched

    def dump(self):
        return self.trie.dump()
