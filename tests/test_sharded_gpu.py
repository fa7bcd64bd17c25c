"""GPU tests of the query-sharded path with the CUDA engine: world 1 (always) and a real 2-rank NCCL
group when the box has >= 2 GPUs (skipped otherwise; the gloo CPU tests cover the host logic)."""
import os
import socket
import sys
from pathlib import Path

import numpy as np
import pytest
import torch

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
sys.path.insert(0, str(ROOT / "tests"))

pytestmark = pytest.mark.gpu


def test_world1_cuda_engine_matches_oracle(cuda_lib, device):
    import oracle
    from _util import assert_forward_exact, assert_grad_close
    from pointcloudregistration_b200 import sharded, synthetic

    b, n, m = 2, 3001, 2500
    x1, x2 = synthetic.make_pair("crosscopy", b, n, m, 70)
    gd1, gd2 = synthetic.grad_pair(b, n, m, 70)
    a = x1.to(device).requires_grad_(True)
    c = x2.to(device).requires_grad_(True)
    d1, d2, i1, i2 = sharded.nnd_sharded_with_idx(a, c)
    ((d1 * gd1.to(device)).sum() + (d2 * gd2.to(device)).sum()).backward()
    want = oracle.forward(x1.numpy(), x2.numpy(), threads=8)
    assert_forward_exact(tuple(t.detach().cpu().numpy() for t in (d1, d2, i1, i2)), want)
    wg = oracle.backward(x1.numpy(), x2.numpy(), gd1.numpy(), gd2.numpy(), want[2], want[3])
    assert_grad_close((a.grad.cpu().numpy(), c.grad.cpu().numpy()), wg)


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _nccl_worker(rank, world, port, b, n, m, out):
    import torch.distributed as dist

    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    torch.cuda.set_device(rank)
    dev = torch.device("cuda", rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)
    try:
        from pointcloudregistration_b200 import sharded, synthetic

        x1, x2 = synthetic.make_pair("duplicates", b, n, m, 71)
        if rank != 0:
            x1, x2 = torch.zeros_like(x1), torch.zeros_like(x2)
        a = x1.to(dev).requires_grad_(True)
        c = x2.to(dev).requires_grad_(True)
        d1, d2, i1, i2 = sharded.nnd_sharded_with_idx(a, c, src=0)
        gd1, gd2 = synthetic.grad_pair(b, n, m, 71)
        ((d1 * gd1.to(dev)).sum() + (d2 * gd2.to(dev)).sum()).backward()
        torch.save({k: v.detach().cpu() for k, v in dict(d1=d1, d2=d2, i1=i1, i2=i2, g1=a.grad, g2=c.grad).items()},
                   f"{out}/r{rank}.pt")
    finally:
        dist.destroy_process_group()


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs 2 GPUs")
def test_two_ranks_nccl(cuda_lib, tmp_path):
    import torch.multiprocessing as mp

    import oracle
    from _util import assert_forward_exact, assert_grad_close
    from pointcloudregistration_b200 import synthetic

    b, n, m = 1, 50001, 40003
    mp.spawn(_nccl_worker, args=(2, _free_port(), b, n, m, str(tmp_path)), nprocs=2, join=True)
    x1, x2 = synthetic.make_pair("duplicates", b, n, m, 71)
    gd1, gd2 = synthetic.grad_pair(b, n, m, 71)
    want = oracle.forward(x1.numpy(), x2.numpy(), threads=16)
    wg = oracle.backward(x1.numpy(), x2.numpy(), gd1.numpy(), gd2.numpy(), want[2], want[3])
    for r in range(2):
        z = torch.load(f"{tmp_path}/r{r}.pt")
        assert_forward_exact(tuple(z[k].numpy() for k in ("d1", "d2", "i1", "i2")), want, what=f"rank {r}: ")
        assert_grad_close((z["g1"].numpy(), z["g2"].numpy()), wg, what=f"rank {r}: ")
