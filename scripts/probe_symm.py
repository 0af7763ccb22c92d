import os, torch, torch.distributed as dist
rank = int(os.environ["RANK"]); local = int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
import torch.distributed._symmetric_memory as symm_mem
t = symm_mem.empty(1 << 20, dtype=torch.uint8, device=torch.device("cuda", local))
hdl = symm_mem.rendezvous(t, dist.group.WORLD.group_name)
print(rank, "ptrs", [hex(p) for p in hdl.buffer_ptrs], "signal", [hex(p) for p in hdl.signal_pad_ptrs][:2], hdl.world_size, flush=True)
t.fill_(rank + 1)
torch.cuda.synchronize(); hdl.barrier()
peer = hdl.get_buffer((rank + 1) % hdl.world_size, (16,), torch.uint8)
print(rank, "peer sees", peer[:4].tolist(), "p2p", torch.cuda.can_device_access_peer(local, (local + 1) % 2), flush=True)
hdl.barrier()
dist.destroy_process_group()
