"""Dev tool (run under torchrun, one rank per GPU): CUDA-event timeline of ONE row-sharded step at C4 -- which phase
of pointcloudregistration_b200/sharded.py costs what, max over ranks (the answer to "where do the non-scan 0.8 ms go").

    python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29517 \
        tools/sharded_timeline.py [n]
"""
import os
import sys
from pathlib import Path

import torch
import torch.distributed as dist

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from pointcloudregistration_b200 import sharded, synthetic  # noqa: E402

local = int(os.environ.get("LOCAL_RANK", "0"))
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
dist.init_process_group("nccl", device_id=dev)
rank, world = dist.get_rank(), dist.get_world_size()
n = int(sys.argv[1]) if len(sys.argv) > 1 else 1_000_000
x1, x2 = (t.to(dev) for t in synthetic.make_pair("uniform", 1, n, n, 4))
gd1, gd2 = (t.to(dev) for t in synthetic.grad_pair(1, n, n, 4, mean=True))


def step():
    d1, d2, i1, i2, info = sharded.sharded_forward(x1, x2, src=0, gather=True)
    return sharded.sharded_backward(x1, x2, gd1, gd2, i1, i2, info, gathered=True)


for _ in range(3):
    step()
rows = {}
for rep in range(5):
    dist.barrier()
    torch.cuda.synchronize()
    sharded.trace = []
    step()
    torch.cuda.synchronize()
    tr, sharded.trace = sharded.trace, None
    for (_, e0), (name, e1) in zip(tr[:-1], tr[1:]):
        rows.setdefault(name, []).append(e0.elapsed_time(e1))
    rows.setdefault("whole step", []).append(tr[0][1].elapsed_time(tr[-1][1]))
names = list(rows)
t = torch.tensor([[min(v) for v in rows.values()], [sum(v) / len(v) for v in rows.values()]], dtype=torch.float64, device=dev)
tmax = t.clone()
dist.all_reduce(tmax, op=dist.ReduceOp.MAX)
if rank == 0:
    print(f"row-sharded step, N=M={n}, {world} ranks; ms per phase on the calling stream, max over ranks (best of 5 | mean of 5)")
    for k, name in enumerate(names):
        print(f"  {name:72s} {tmax[0, k].item():9.3f} | {tmax[1, k].item():9.3f}")
dist.destroy_process_group()
