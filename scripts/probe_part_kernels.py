"""Debug helper: per-kernel-class times of the three phases of the partitioned router
(torchrun --nproc-per-node N scripts/probe_part_kernels.py)."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "vllm-production-stack_b200"))
import numpy as np, torch, torch.distributed as dist
from b200_router import _lib, traces
from b200_router.distributed import PartitionedPrefixRouter
from b200_router.packing import pack_arrays
rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local); dev = torch.device("cuda", local)
dist.init_process_group("nccl", device_id=dev)
nb, per = 9, 65536
tr = traces.SharedPrefixTrace(n_total=nb * world * per, batch=per, n_sys=64, sys_chars=2048, suffix_lo=256, suffix_hi=1024, seed=20240607)
res = [pack_arrays(*tr.batch_arrays(b * world + rank), pin=False).to(dev) for b in range(nb)]
mc = max(p.max_chunks for p in res)
r = PartitionedPrefixRouter(cap_log2=int(os.environ.get("CAPLOG2", "25")), device=dev, max_requests=per, max_chunks=mc)
out = (torch.empty(per, dtype=torch.int32, device=dev), torch.empty(per, dtype=torch.int32, device=dev), torch.empty(per, dtype=torch.int64, device=dev))
for b in range(3):
    r.route(res[b], 0xFF, 1, 0, out=out)
dist.barrier(); torch.cuda.synchronize()
_lib.prof_enable(True)
acc = {}
for b in range(3, nb):
    for name, fn in (("send", lambda: r.phase_send(res[b])), ("route", lambda: r.phase_route(0xFF, 1)),
                     ("finish", lambda: r.phase_finish(out))):
        fn()
        torch.cuda.synchronize()
        for k, v in _lib.prof_read().items():
            if v[1]:
                acc.setdefault((name, k), []).append(v[0] * 1e3)
    dist.barrier()
_lib.prof_enable(False)
print("rank %d: " % rank + ", ".join("%s/%s %.1f" % (a, k, float(np.mean(v))) for (a, k), v in acc.items()), flush=True)
dist.destroy_process_group()
