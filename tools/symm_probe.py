import os, torch, torch.distributed as dist
rank=int(os.environ["RANK"]); local=int(os.environ["LOCAL_RANK"]); world=int(os.environ["WORLD_SIZE"])
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
import torch.distributed._symmetric_memory as symm
t = symm.empty(1<<20, dtype=torch.uint8, device=torch.device("cuda", local))
h = symm.rendezvous(t, dist.group.WORLD.group_name)
print(rank, "ptrs", [hex(p) for p in h.buffer_ptrs], "signal", [hex(p) for p in h.signal_pad_ptrs][:2], flush=True)
t.fill_(rank+1)
dist.barrier(); torch.cuda.synchronize()
# read peer memory through the mapped pointer using a tiny torch op: get_buffer
peer = h.get_buffer((rank+1)%world, (16,), torch.uint8)
print(rank, "peer first bytes", peer[:4].tolist(), flush=True)
dist.destroy_process_group()
