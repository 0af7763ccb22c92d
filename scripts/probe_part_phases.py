"""Debug helper: per-phase CUDA-event timing of the partitioned router, unsynchronised steps."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "vllm-production-stack_b200"))
import numpy as np, torch, torch.distributed as dist
from b200_router import traces
from b200_router.distributed import PartitionedPrefixRouter
from b200_router.packing import pack_arrays
rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local); dev = torch.device("cuda", local)
dist.init_process_group("nccl", device_id=dev)
nb, per = 11, 65536
tr = traces.SharedPrefixTrace(n_total=nb * world * per, batch=per, n_sys=64, sys_chars=2048, suffix_lo=256, suffix_hi=1024, seed=20240607)
res = [pack_arrays(*tr.batch_arrays(b * world + rank), pin=False).to(dev) for b in range(nb)]
mc = max(p.max_chunks for p in res)
r = PartitionedPrefixRouter(cap_log2=int(os.environ.get("CAPLOG2", "25")), device=dev, max_requests=per, max_chunks=mc)
out = (torch.empty(per, dtype=torch.int32, device=dev), torch.empty(per, dtype=torch.int32, device=dev), torch.empty(per, dtype=torch.int64, device=dev))
for b in range(3):
    r.route(res[b], 0xFF, 1, 0, out=out)
dist.barrier(); torch.cuda.synchronize()
ev = []
for b in range(3, nb):
    e = [torch.cuda.Event(enable_timing=True) for _ in range(4)]
    e[0].record(); r.phase_send(res[b]); e[1].record(); r.phase_route(0xFF, 1); e[2].record(); r.phase_finish(out); e[3].record()
    ev.append(e)
torch.cuda.synchronize()
tot = ev[0][0].elapsed_time(ev[-1][3]) / len(ev)
ph = np.array([[e[i].elapsed_time(e[i + 1]) for i in range(3)] for e in ev]) * 1e3
print("rank %d: %.0f us/step; send %s route %s finish %s" % (rank, tot * 1e3, ph[:, 0].round().tolist(), ph[:, 1].round().tolist(), ph[:, 2].round().tolist()), flush=True)
dist.destroy_process_group()
