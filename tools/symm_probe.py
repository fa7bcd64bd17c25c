"""Dev probe (2+ GPUs): does torch's symmetric memory work on this box?  torchrun --nproc-per-node N tools/symm_probe.py"""
import os

import torch
import torch.distributed as dist
import torch.distributed._symmetric_memory as symm_mem

rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
dist.init_process_group("nccl", device_id=dev)
t = symm_mem.empty(1 << 20, dtype=torch.int64, device=dev)
hdl = symm_mem.rendezvous(t, dist.group.WORLD)
t.fill_(rank + 1)
hdl.barrier(channel=0)
peer = hdl.get_buffer((rank + 1) % world, (1 << 20,), torch.int64)
val = int(peer[12345].item())
print(f"rank {rank}: peer value {val} (want {(rank + 1) % world + 1}); buffer_ptrs {len(hdl.buffer_ptrs)}, "
      f"multicast {hdl.has_multicast_support(dev.type, dev.index) if hasattr(hdl, 'has_multicast_support') else '?'}, "
      f"signal pad {hdl.signal_pad_size}", flush=True)
hdl.barrier(channel=0)
# timing of a barrier
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(100):
    hdl.barrier(channel=0)
e1.record()
torch.cuda.synchronize()
if rank == 0:
    print(f"barrier: {e0.elapsed_time(e1) * 10:.2f} us each", flush=True)
dist.destroy_process_group()
