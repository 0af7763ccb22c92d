"""The partitioned exchange with world = 1 on ONE GPU (every request is owned by this rank):
all of its kernels run -- place, bulk, wait, gather, claim, select, commit, push, wait,
scatter -- so that a launch list (ncu) shows each one alone.  Usage:
    python scripts/profile_part_single.py [batches] [batch]"""
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "vllm-production-stack_b200"))
import torch  # noqa: E402
from b200_router import traces  # noqa: E402
from b200_router.distributed import PartitionedPrefixRouter  # noqa: E402
from b200_router.packing import pack_arrays  # noqa: E402

nb = int(sys.argv[1]) if len(sys.argv) > 1 else 6
batch = int(sys.argv[2]) if len(sys.argv) > 2 else 65536
dev = torch.device("cuda:0")
tr = traces.SharedPrefixTrace(n_total=nb * batch, batch=batch, n_sys=64, sys_chars=2048,
                              suffix_lo=256, suffix_hi=1024, seed=1234)
packs = [pack_arrays(*tr.batch_arrays(b), pin=False).to(dev) for b in range(nb)]
mr, mc = max(p.n for p in packs), max(p.max_chunks for p in packs)
nbytes = PartitionedPrefixRouter.symm_bytes_for(1, mr, mc)
buf = torch.zeros(nbytes, dtype=torch.uint8, device=dev)
r = PartitionedPrefixRouter(cap_log2=24, device=dev, max_requests=mr, max_chunks=mc,
                            _sim={"world": 1, "rank": 0, "buffers": [buf]})
out = (torch.empty(mr, dtype=torch.int32, device=dev), torch.empty(mr, dtype=torch.int32, device=dev),
       torch.empty(mr, dtype=torch.int64, device=dev))
# time spent inside the native call, to separate it from the Python wrapper around it
from b200_router import _lib  # noqa: E402
_L = _lib.lib()
_native = _L.b200r_part_step
_acc = [0.0, 0]


def _timed(*a):
    t = time.perf_counter()
    rc = _native(*a)
    _acc[0] += time.perf_counter() - t
    _acc[1] += 1
    return rc


_L.b200r_part_step = _timed
torch.cuda.synchronize()
warm = min(4, nb - 1)
for p in packs[:warm]:
    r.route(p, 0xFF, 7, out=out, inputs_ready=True)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
t0 = time.perf_counter()
e0.record()
_acc[0], _acc[1] = 0.0, 0
for p in packs[warm:]:
    r.route(p, 0xFF, 7, out=out, inputs_ready=True)
host = (time.perf_counter() - t0) / (nb - warm)
e1.record()
torch.cuda.synchronize()
print("batches %d x %d (after %d warm-up batches): %.1f us per batch on the GPU, %.1f us of host time per call "
      "(%.1f us inside b200r_part_step)" %
      (nb - warm, batch, warm, e0.elapsed_time(e1) * 1e3 / (nb - warm), host * 1e6, _acc[0] / max(_acc[1], 1) * 1e6),
      r.check_status())
