"""Time the chunk-hashing stage alone on a C3-shaped batch (CUDA events, distinct batches so that
inputs are not in L2):  python scripts/microbench/hash_time.py [hash_mode ...]"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(ROOT, "vllm-production-stack_b200"))
import torch  # noqa: E402
from b200_router import _lib, traces  # noqa: E402
from b200_router.packing import pack_arrays  # noqa: E402
from b200_router.trie import GpuHashTrie  # noqa: E402

modes = [int(x) for x in sys.argv[1:]] or [1, 0]
dev = torch.device("cuda:0")
nb, batch = 10, 65536
tr = traces.SharedPrefixTrace(n_total=nb * batch, batch=batch, n_sys=64, sys_chars=2048,
                              suffix_lo=256, suffix_hi=1024, seed=1234)
packs = [pack_arrays(*tr.batch_arrays(b), pin=False).to(dev) for b in range(nb)]
t = GpuHashTrie(cap_log2=16, device=dev)
payload = sum(p.payload_bytes for p in packs[2:])
for m in modes:
    _lib.check(_lib.lib().b200r_set_hash_mode(m))
    _lib.prof_enable(True)
    for p in packs[:2]:
        t.hash_chunks(p)
    torch.cuda.synchronize()
    _lib.prof_read()
    for p in packs[2:]:
        t.hash_chunks(p)
    torch.cuda.synchronize()
    prof = _lib.prof_read()
    _lib.prof_enable(False)
    ms = prof["hash"][0] / max(prof["hash"][1], 1)
    print("mode %d: hash kernel %.1f us per batch, %.0f GB/s; prep %.1f us" %
          (m, ms * 1e3, payload / (nb - 2) / (ms * 1e-3) / 1e9, prof["prep"][0] / (nb - 2) * 1e3))
