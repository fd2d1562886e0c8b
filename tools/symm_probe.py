import os, torch, torch.distributed as dist
rank = int(os.environ["RANK"]); world = int(os.environ["WORLD_SIZE"]); local = int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local); dev = torch.device("cuda", local)
dist.init_process_group("nccl", device_id=dev)
import torch.distributed._symmetric_memory as symm_mem
t = symm_mem.empty(1 << 16, dtype=torch.uint8, device=dev)
h = symm_mem.rendezvous(t, dist.group.WORLD)
if rank == 0:
    print("handle type", type(h).__name__)
    print("attrs", [a for a in dir(h) if not a.startswith("_")])
    print("buffer_ptrs", [hex(p) for p in h.buffer_ptrs], "signal_pad_ptrs", [hex(p) for p in h.signal_pad_ptrs])
    print("buffer_size", h.buffer_size, "signal_pad_size", getattr(h, "signal_pad_size", None), "rank", h.rank, "world", h.world_size)
t.fill_(rank + 1)
h.barrier()
peer = h.get_buffer((rank + 1) % world, (16,), torch.uint8)
print(rank, "peer view first bytes", peer[:4].tolist(), flush=True)
h.barrier()
dist.destroy_process_group()
