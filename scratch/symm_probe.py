import os, torch, torch.distributed as dist
rank = int(os.environ["RANK"]); world = int(os.environ["WORLD_SIZE"])
torch.cuda.set_device(int(os.environ["LOCAL_RANK"]))
dist.init_process_group("nccl", device_id=torch.device("cuda", int(os.environ["LOCAL_RANK"])))
import torch.distributed._symmetric_memory as symm
t = symm.empty(1 << 20, dtype=torch.float32, device=torch.device("cuda", int(os.environ["LOCAL_RANK"])))
h = symm.rendezvous(t, group=dist.group.WORLD)
print(rank, "ptrs", [hex(p) for p in h.buffer_ptrs], "rank", h.rank, "world", h.world_size, type(h).__name__, flush=True)
t.fill_(float(rank))
h.barrier()
peer = h.get_buffer((rank + 1) % world, (4,), torch.float32)
print(rank, "peer value", peer.tolist(), flush=True)
h.barrier()
dist.destroy_process_group()
