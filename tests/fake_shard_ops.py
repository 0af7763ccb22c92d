"""CPU emulation of the two device steps of the replicated multi-GPU design (hash_chunks /
route_hashes) with the same tensor interfaces as b200_router.distributed.CudaShardOps.

TEST INFRASTRUCTURE: lets the multi-rank orchestration (ShardedPrefixRouter +
torch.distributed, gloo backend, world_size 2) run on machines without a GPU.  Chunk hashes
come from the oracle (oracle.c); routing from hashes is the sequential reference semantics
(hashtrie.py:58-103, routing_logic.py:414-423) over a dict keyed by hash path.
"""
import numpy as np
import torch

from oracle import coracle
from oracle.chooser import pick

M64 = (1 << 64) - 1


def _s64(x):
    return x - (1 << 64) if x >= (1 << 63) else x


class FakeShardOps:
    def __init__(self, chunk_chars=128):
        self.device = torch.device("cpu")
        self.nodes = {}          # hash path (tuple) -> mask
        self.root_mask = 0
        self.cc = chunk_chars

    def hash_chunks(self, p):
        buf = p.bytes.numpy()
        base, hs = [0], []
        for i in range(p.n):
            s, l = int(p.starts[i]), int(p.lens[i])
            h = coracle.xxh64_chunks(buf[s:s + l].tobytes(), self.cc)
            hs += [_s64(int(x)) for x in h]
            base.append(len(hs))
        return (torch.tensor(base, dtype=torch.int32),
                torch.tensor(hs + [0], dtype=torch.int64))

    def route_hashes(self, hashes, chunk_base, n, max_chunks, avail, seed, seq0):
        hs = hashes.numpy().view(np.uint64)
        cb = chunk_base.numpy()
        ep = np.empty(n, dtype=np.int32)
        depth = np.empty(n, dtype=np.int32)
        matched = np.empty(n, dtype=np.uint64)
        for i in range(n):
            path = tuple(int(x) for x in hs[cb[i]:cb[i + 1]])
            sel, d = avail, 0
            for k in range(1, len(path) + 1):
                m = self.nodes.get(path[:k])
                if m is None or not (m & sel):
                    break
                sel, d = m & sel, k
            cnt = bin(sel).count("1")
            if cnt:
                k = pick(seed, seq0 + i, cnt)
                e = [b for b in range(64) if sel >> b & 1][k]
                self.root_mask |= 1 << e
                for k2 in range(1, len(path) + 1):
                    self.nodes[path[:k2]] = self.nodes.get(path[:k2], 0) | (1 << e)
            else:
                e = -1
            ep[i], depth[i], matched[i] = e, d, sel
        return (torch.from_numpy(ep), torch.from_numpy(depth),
                torch.from_numpy(matched.view(np.int64)))

    def check_status(self, count_nodes=False):
        return {"status": 0, "num_nodes": len(self.nodes)}

    def dump(self):
        out = dict(self.nodes)
        out[()] = self.root_mask
        return out
