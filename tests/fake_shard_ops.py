"""CPU emulation of the four per-batch shard steps (classify / apply / select /
commit) with the same tensor interfaces as b200_router.distributed.CudaShardOps.

TEST INFRASTRUCTURE: lets the multi-rank orchestration (ShardedPrefixRouter +
torch.distributed, gloo backend, world_size 2) run on machines without a GPU.
Chunk hashes come from the oracle (oracle.c); path keys use the library's own
key recurrence constants so that dumps decode with b200_router.trie.decode_dump.
"""
import numpy as np
import torch

from b200_router import _lib
from b200_router.trie import decode_dump, key_constants
from oracle import coracle
from oracle.chooser import pick

M64 = (1 << 64) - 1


def _s64(x):
    return x - (1 << 64) if x >= (1 << 63) else x


def _u64(x):
    return int(x) & M64


class FakeShardOps:
    def __init__(self, chunk_chars=128):
        self.device = torch.device("cpu")
        self.a1, self.a2, self.r1, self.r2, _, _ = key_constants()
        self.table = {}          # (k1,k2) -> [mask, first_epoch, first_req, h]
        self.root_mask = 0
        self.epoch = 0
        self.cc = chunk_chars

    def _keys(self, hashes):
        k1, k2, out = self.r1, self.r2, []
        for h in hashes:
            h = int(h)
            k1 = (self.a1 * k1 + h) & M64
            k2 = (self.a2 * k2 + (((h << 32) | (h >> 32)) & M64)) & M64
            out.append((k1 or 1, k2 or 1))
        return out

    def classify(self, p, avail, recs, count):
        self.epoch += 1
        buf = p.bytes.numpy()
        self.chunks = []         # per request: list of (key, h)
        self.preok = []
        rows = []
        for i in range(p.n):
            s, l = int(p.starts[i]), int(p.lens[i])
            hs = coracle.xxh64_chunks(buf[s:s + l].tobytes(), self.cc)
            keys = self._keys(hs)
            self.chunks.append(list(zip(keys, [int(h) for h in hs])))
            flags = []
            for d, (k, h) in enumerate(self.chunks[-1]):
                ent = self.table.get(k)
                ok = ent is not None and (ent[0] & avail) != 0
                flags.append(ok)
                if not ok:
                    rows.append([_s64(k[0]), _s64(k[1]), _s64(h), _s64(i | (d << 32))])
            self.preok.append(flags)
        if rows:
            recs[:len(rows)] = torch.tensor(rows, dtype=torch.int64)
        count[0] = len(rows)
        count[1] = p.n

    def apply(self, recs_all, meta_all, world, stride, rank, rec_slot):
        ra = recs_all.numpy()
        offs = np.concatenate([[0], np.cumsum(meta_all[:, 1].numpy())])
        for q in range(world):
            for j in range(int(meta_all[q, 0])):
                k1, k2, h, rd = (_u64(x) for x in ra[q, j])
                req = (rd & 0xFFFFFFFF) + int(offs[q])
                ent = self.table.setdefault((k1, k2), [0, 0, 0, h])
                if ent[1] != self.epoch or req < ent[2]:
                    ent[1], ent[2] = self.epoch, req

    def select(self, avail, seed, seq0, my_off, leader, choice, depth, matched):
        for i, ch in enumerate(self.chunks):
            dstar, star = 0, None
            for d, (k, h) in enumerate(ch):
                ent = self.table.get(k)
                if self.preok[i][d]:
                    ok = True
                else:
                    ok = ent is not None and ent[1] == self.epoch and ent[2] < my_off + i
                if not ok:
                    break
                dstar, star = d + 1, (k, self.preok[i][d])
            depth[i] = dstar
            if dstar == 0 or star[1]:
                cand = avail if dstar == 0 else (self.table[star[0]][0] & avail)
                bits = [b for b in range(64) if (cand >> b) & 1]
                leader[i] = -1
                choice[i] = bits[pick(seed, seq0 + my_off + i, len(bits))] if bits else -1
                matched[i] = _s64(cand)
            else:
                leader[i] = self.table[star[0]][2]
                choice[i] = -1

    def commit(self, recs_all, meta_all, world, stride, rec_slot, lcd_all, rstride, rank,
               endpoint_all, endpoint, depth, matched):
        offs = [0] + [int(x) for x in np.cumsum(meta_all[:, 1].numpy())]
        counts = meta_all[:, 0]
        leader_all = lcd_all[:, 0, :].contiguous()
        choice_all = lcd_all[:, 1, :].contiguous()
        depth_all = lcd_all[:, 2, :].contiguous()

        def index_of(pos):
            q = 0
            while q + 1 < world and pos >= offs[q + 1]:
                q += 1
            return q * rstride + (pos - offs[q])
        la, ca, da = leader_all.reshape(-1), choice_all.reshape(-1), depth_all.reshape(-1)
        ep_all = {}
        for q in range(world):
            for i in range(offs[q + 1] - offs[q]):
                j = q * rstride + i
                while int(la[j]) >= 0:
                    j = index_of(int(la[j]))
                e = int(ca[j])
                ep_all[q * rstride + i] = e
                if e >= 0:
                    self.root_mask |= 1 << e
        ra = recs_all.numpy()
        for q in range(world):
            for j in range(int(counts[q])):
                k1, k2, h, rd = (_u64(x) for x in ra[q, j])
                req, d = rd & 0xFFFFFFFF, rd >> 32
                e = ep_all[q * rstride + req]
                if e >= 0 and d >= int(da[q * rstride + req]):
                    self.table[(k1, k2)][0] |= 1 << e
        n = offs[rank + 1] - offs[rank]
        for i in range(n):
            e = ep_all[rank * rstride + i]
            endpoint[i] = e
            depth[i] = int(da[rank * rstride + i])
            if int(la[rank * rstride + i]) >= 0:
                matched[i] = _s64(1 << e) if e >= 0 else 0

    def check_status(self, count_nodes=False):
        return {"status": 0, "num_nodes": len(self.table), "capacity": 0, "root_mask": self.root_mask}

    def dump(self):
        arr = np.array([[k[0], k[1], v[3], v[0]] for k, v in self.table.items()],
                       dtype=np.uint64).reshape(-1, 4)
        return decode_dump(arr, self.root_mask)
