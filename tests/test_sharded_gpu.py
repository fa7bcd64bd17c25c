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
    # deterministic mode of the sharded path (replicated nnd_backward_exact_cuda_f32): every gradient bit
    a2 = x1.to(device).requires_grad_(True)
    c2 = x2.to(device).requires_grad_(True)
    e1, e2, _, _ = sharded.nnd_sharded_with_idx(a2, c2, exact=True)
    ((e1 * gd1.to(device)).sum() + (e2 * gd2.to(device)).sum()).backward()
    assert a2.grad.cpu().numpy().tobytes() == wg[0].tobytes() and c2.grad.cpu().numpy().tobytes() == wg[1].tobytes()


@pytest.mark.parametrize("kind,b,n,m,cuts", [
    ("crosscopy", 1, 5000, 4100, (0, 2048, 5000)),
    ("lattice", 2, 3001, 2500, (0, 8, 1000, 1000, 3001)),   # includes an empty range and a one-group range
    ("identical", 1, 4096, 4096, (0, 4088, 4096)),
    ("duplicates", 3, 700, 1300, (0, 696, 700)),
])
def test_row_ranges_merge_to_the_unsharded_result(cuda_lib, device, kind, b, n, m, cuts):
    """nnd_forward_rows over disjoint row ranges + integer MIN of the column keys + column finalize
    == the oracle, bit for bit (what the multi-GPU path does, on one GPU)."""
    import oracle
    from _util import assert_forward_exact
    from pointcloudregistration_b200 import functional as F
    from pointcloudregistration_b200 import synthetic

    x1, x2 = synthetic.make_pair(kind, b, n, m, 72)
    a, c = x1.to(device), x2.to(device)
    d1 = torch.empty(b, n, device=device)
    i1 = torch.empty(b, n, dtype=torch.int32, device=device)
    merged = None
    for lo, hi in zip(cuts[:-1], cuts[1:]):
        dr, ir, keys = F.nnd_forward_rows_raw(a, c, lo, hi)
        d1[:, lo:hi], i1[:, lo:hi] = dr, ir
        merged = keys if merged is None else torch.minimum(merged, keys)
    d2, i2 = F.nnd_finalize_cols_raw(a, c, merged)
    want = oracle.forward(x1.numpy(), x2.numpy(), threads=8)
    assert_forward_exact(tuple(t.cpu().numpy() for t in (d1, d2, i1, i2)), want)


@pytest.mark.parametrize("kind", ["uniform", "lattice"])
def test_row_ranges_at_scale_equal_the_single_launch(cuda_lib, device, kind):
    """Size-independent property at N=M=262144 (symmetric kernel, large shape): 3 uneven row ranges merged by integer
    MIN give bit-identical dist/idx to the single-launch forward (which the C5/C4 tests tie to the oracle)."""
    from pointcloudregistration_b200 import functional as F
    from pointcloudregistration_b200 import synthetic

    n = m = 262144
    x1, x2 = synthetic.make_pair(kind, 1, n, m, 73)
    a, c = x1.to(device), x2.to(device)
    want = F.nnd_forward_raw(a, c)
    d1 = torch.empty(1, n, device=device)
    i1 = torch.empty(1, n, dtype=torch.int32, device=device)
    merged = None
    for lo, hi in ((0, 100000), (100000, 100008), (100008, n)):
        dr, ir, keys = F.nnd_forward_rows_raw(a, c, lo, hi)
        d1[:, lo:hi], i1[:, lo:hi] = dr, ir
        merged = keys if merged is None else torch.minimum(merged, keys)
    d2, i2 = F.nnd_finalize_cols_raw(a, c, merged)
    for got, w in zip((d1, d2, i1, i2), want):
        assert torch.equal(got, w)


@pytest.mark.parametrize("kind,b,n,m,cuts", [
    ("crosscopy", 1, 5000, 4100, (0, 2048, 5000)),             # second range longer than M: normal orientation
    ("lattice", 2, 3001, 2500, (0, 8, 1000, 1000, 3001)),      # empty range, one-group range
    ("identical", 1, 4096, 9000, (0, 4088, 4096)),
    ("duplicates", 3, 700, 1300, (0, 696, 700)),
    ("clustered", 1, 20000, 12000, (0, 4000, 4008, 11000, 20000)),
    ("uniform", 2, 9000, 30000, (0, 3000, 9000)),              # every range shorter than M: role-swapped
])
def test_row_ranges_through_the_panel_scans(cuda_lib, device, kind, b, n, m, cuts):
    """Same as above with every scan forced through the panel kernel: ranges shorter than M run role-swapped (xyz2 in
    the registers, the range streamed; its partial results come back as column keys), the others in the normal
    orientation.  Merged result == oracle, bit for bit."""
    import oracle
    from _util import assert_forward_exact
    from pointcloudregistration_b200 import functional as F
    from pointcloudregistration_b200 import synthetic

    x1, x2 = synthetic.make_pair(kind, b, n, m, 74)
    a, c = x1.to(device), x2.to(device)
    d1 = torch.full((b, n), -1.0, device=device)
    i1 = torch.full((b, n), -1, dtype=torch.int32, device=device)
    before = cuda_lib.nnd_set_panel_min_rows(1)
    try:
        merged = None
        for lo, hi in zip(cuts[:-1], cuts[1:]):
            dr, ir, keys = F.nnd_forward_rows_raw(a, c, lo, hi)
            d1[:, lo:hi], i1[:, lo:hi] = dr, ir
            merged = keys if merged is None else torch.minimum(merged, keys)
        d2, i2 = F.nnd_finalize_cols_raw(a, c, merged)
    finally:
        cuda_lib.nnd_set_panel_min_rows(before)
    want = oracle.forward(x1.numpy(), x2.numpy(), threads=8)
    assert_forward_exact(tuple(t.cpu().numpy() for t in (d1, d2, i1, i2)), want)


def test_swapped_row_range_with_nonfinite_points(cuda_lib, device):
    """NaN-first rule across role-swapped ranges: the rule is applied once, after the merge, not per range."""
    import numpy as np
    import oracle
    from pointcloudregistration_b200 import functional as F
    from pointcloudregistration_b200 import synthetic

    x1 = synthetic.uniform(1, 4000, 75).numpy()
    x2 = synthetic.uniform(1, 9000, 76).numpy()
    x1[0, 0, 1] = np.nan      # first row of xyz1: every column's first candidate is NaN
    x1[0, 2500] = np.inf
    x2[0, 0] = np.nan         # first column: every row's first candidate is NaN
    x2[0, 77, 2] = np.nan
    a, c = torch.from_numpy(x1).to(device), torch.from_numpy(x2).to(device)
    d1 = torch.empty(1, 4000, device=device)
    i1 = torch.empty(1, 4000, dtype=torch.int32, device=device)
    before = cuda_lib.nnd_set_panel_min_rows(1)
    try:
        merged = None
        for lo, hi in ((0, 1600), (1600, 4000)):
            dr, ir, keys = F.nnd_forward_rows_raw(a, c, lo, hi)
            d1[:, lo:hi], i1[:, lo:hi] = dr, ir
            merged = keys if merged is None else torch.minimum(merged, keys)
        d2, i2 = F.nnd_finalize_cols_raw(a, c, merged)
    finally:
        cuda_lib.nnd_set_panel_min_rows(before)
    w1, w2, j1, j2 = oracle.forward(x1, x2)
    assert np.array_equal(i1.cpu().numpy(), j1) and np.array_equal(i2.cpu().numpy(), j2)
    for got, want in ((d1, w1), (d2, w2)):
        g = got.cpu().numpy()
        assert np.array_equal(np.isnan(g), np.isnan(want)) and g[~np.isnan(g)].tobytes() == want[~np.isnan(want)].tobytes()


def test_row_range_must_be_whole_groups(cuda_lib, device):
    from pointcloudregistration_b200 import functional as F

    a = torch.rand(1, 100, 3, device=device)
    c = torch.rand(1, 50, 3, device=device)
    g = F.row_group()
    with pytest.raises(RuntimeError):
        F.nnd_forward_rows_raw(a, c, 1, 100)       # start inside a group
    with pytest.raises(RuntimeError):
        F.nnd_forward_rows_raw(a, c, 0, g + 1)     # end inside a group (and not at N)
    F.nnd_forward_rows_raw(a, c, g, 100)           # ends at N: fine


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
        used_peer_memory = len(sharded.PeerExchange._cache) > 0
        # the same step through the NCCL collectives: the forward must agree bit for bit, the gradients to rounding
        sharded.peer_exchange = False
        a2 = a.detach().clone().requires_grad_(True)
        c2 = c.detach().clone().requires_grad_(True)
        e1, e2, j1, j2 = sharded.nnd_sharded_with_idx(a2, c2, src=0)
        ((e1 * gd1.to(dev)).sum() + (e2 * gd2.to(dev)).sum()).backward()
        sharded.peer_exchange = True
        same_forward = all(bool(torch.equal(p_, q_)) for p_, q_ in ((d1, e1), (d2, e2), (i1, j1), (i2, j2)))
        scale = float(max(a2.grad.abs().max(), c2.grad.abs().max()))
        grad_gap = float(max((a.grad - a2.grad).abs().max(), (c.grad - c2.grad).abs().max())) / scale
        torch.save(dict(d1=d1.detach().cpu(), d2=d2.detach().cpu(), i1=i1.cpu(), i2=i2.cpu(), g1=a.grad.cpu(), g2=c.grad.cpu(),
                        used_peer_memory=used_peer_memory, same_forward=same_forward, grad_gap=grad_gap),
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
        # the exchange over peer memory (when this box has symmetric memory) and the NCCL collectives agree
        assert z["same_forward"] and z["grad_gap"] <= 1e-5, (r, z["used_peer_memory"], z["grad_gap"])
    z0, z1 = torch.load(f"{tmp_path}/r0.pt"), torch.load(f"{tmp_path}/r1.pt")
    if z0["used_peer_memory"]:  # the two-shot sum runs in rank order: the same gradient bits on every rank
        assert torch.equal(z0["g1"], z1["g1"]) and torch.equal(z0["g2"], z1["g2"])
