import sys, os
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "vllm-production-stack_b200"))
import numpy as np, torch
from b200_router import traces
from b200_router.packing import pack_arrays
from b200_router.trie import GpuHashTrie
from oracle import coracle
tr = traces.SharedPrefixTrace(n_total=1 << 16, batch=1 << 16, n_sys=64, sys_chars=2048, suffix_lo=256, suffix_hi=1024, seed=7)
buf, starts, lens = tr.batch_arrays(0)
orc = coracle.TrieOracle()
oep, odepth, om = orc.route_batch(buf, starts, lens, 0xFF, 99, 0)
p = pack_arrays(buf, starts, lens, pin=False)
import ctypes
from b200_router import _lib
mode=int(sys.argv[1]) if len(sys.argv)>1 else 0
nbad = 0
for it in range(40):
    t = GpuHashTrie(cap_log2=21 if it % 2 else 24)
    ep, depth, matched = t.route(p, 0xFF, 99, 0)
    e, d = ep.cpu().numpy(), depth.cpu().numpy()
    bad = np.nonzero((e != oep) | (d != odepth))[0]
    st = t.status()
    if len(bad) or st["num_nodes"] != orc.num_nodes() - 1 or st["status"]:
        nbad += 1
        arr, root = t.dump_raw()
        idx = (arr[:, 0] & np.uint64(0xFFFFFFFF)).astype(np.int64)
        rem = (arr[:, 0] >> np.uint64(32)).astype(np.int64)
        par = (arr[:, 1] & np.uint64(0xFFFFFFFF)).astype(np.int64)
        head = (arr[:, 1] >> np.uint64(63)).astype(np.int64)
        key = par.astype(np.uint64) * np.uint64(0x9E3779B97F4A7C15) ^ arr[:, 2]
        u, c = np.unique(key, return_counts=True)
        dups = u[c > 1]
        print("iter", it, "bad", len(bad), bad[:6], "depth", d[bad[:6]], odepth[bad[:6]], "nodes", st["num_nodes"], orc.num_nodes() - 1, "status", st["status"], "dupkeys", len(dups))
        for k in dups[:3]:
            rows = np.nonzero(key == k)[0]
            for r_ in rows:
                print("   dup node idx", idx[r_], "rem", rem[r_], "parent", par[r_], "head", head[r_], "h %016x" % int(arr[r_, 2]), "mask", int(arr[r_, 3]))
        # unique idx?
        ui, ci = np.unique(idx, return_counts=True)
        print("   records listed twice:", int((ci > 1).sum()))
    t.close()
print("mode", mode, "bad iterations", nbad, "of 40")
