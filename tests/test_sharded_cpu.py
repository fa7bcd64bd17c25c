"""CPU tests of the multi-GPU host logic: world_size-2 (and 3) `gloo` process groups with the oracle
injected as the per-rank engine.  Checks: forward bit-identical to the unsharded oracle for any G,
gradients within 1e-5, broadcast from src, uneven shards."""
import os
import socket
import sys
from pathlib import Path

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
sys.path.insert(0, str(ROOT / "tests"))


KEY_INIT = np.int64(0x7fffffffffffffff)
TT = 512  # column keys are padded to whole 512-point tiles (include/nnd_b200.h: nnd_colkeys_count)
GROUP = 8


def _dense_sqdist(p1, p2):
    """D[j,k] exactly as my_lib.cpp:12-15 evaluates it for query p1[j], target p2[k] (numpy f32 ops are
    separately rounded IEEE operations, no FMA)."""
    d = p2[None, :, :] - p1[:, None, :]
    sq = d * d
    return (sq[..., 0] + sq[..., 1]) + sq[..., 2]


class OracleCompute:
    """Test-only engine: same interface as sharded.CudaCompute, backed by oracle/ and numpy (CPU)."""

    @staticmethod
    def row_group():
        return GROUP

    @staticmethod
    def rows(xyz1, xyz2, lo, hi):
        import oracle

        assert (lo % GROUP == 0 or lo == xyz1.size(1)) and (hi % GROUP == 0 or hi == xyz1.size(1))
        b, m = xyz1.size(0), xyz2.size(1)
        per = (m + TT - 1) // TT * TT
        keys = np.full((b, per), KEY_INIT, np.int64)
        if hi > lo:
            d, i = oracle.query(xyz1[:, lo:hi].contiguous().numpy(), xyz2.numpy())
            for bi in range(b):
                D = _dense_sqdist(xyz1[bi, lo:hi].numpy(), xyz2[bi].numpy())  # [rows, m]
                first = D.argmin(axis=0)  # lowest row attaining the column minimum
                bits = D[first, np.arange(m)].view(np.int32).astype(np.int64)
                keys[bi, :m] = (bits << 32) | ((first + lo) // GROUP)
        else:
            d = np.zeros((b, 0), np.float32)
            i = np.zeros((b, 0), np.int32)
        return torch.from_numpy(d), torch.from_numpy(i), torch.from_numpy(keys.reshape(-1))

    @staticmethod
    def finalize_cols(xyz1, xyz2, colkeys):
        b, n, m = xyz1.size(0), xyz1.size(1), xyz2.size(1)
        keys = colkeys.numpy().reshape(b, -1)[:, :m]
        dist = (keys >> 32).astype(np.int32).view(np.float32)
        group = (keys & 0xffffffff).astype(np.int64)
        idx = np.zeros((b, m), np.int32)
        for bi in range(b):
            D = _dense_sqdist(xyz1[bi].numpy(), xyz2[bi].numpy())
            for k in range(m):
                rows = np.arange(group[bi, k] * GROUP, min(group[bi, k] * GROUP + GROUP, n))
                hit = rows[D[rows, k] == dist[bi, k]]
                idx[bi, k] = hit[0]
        return torch.from_numpy(np.ascontiguousarray(dist)), torch.from_numpy(idx)

    @staticmethod
    def backward_term(q, t, gd, idx, grad_q, grad_t):
        qn, tn, g, ix = q.numpy(), t.numpy(), gd.numpy(), idx.numpy().astype(np.int64)
        for b in range(qn.shape[0]):
            tt = (g[b] * np.float32(2))[:, None] * (qn[b] - tn[b][ix[b]])
            grad_q[b] += torch.from_numpy(tt)
            np.subtract.at(grad_t[b].numpy(), ix[b], tt)


    @staticmethod
    def backward_exact(xyz1, xyz2, gd1, gd2, idx1, idx2):
        import oracle

        g1, g2 = oracle.backward(xyz1.detach().numpy(), xyz2.detach().numpy(), gd1.numpy(), gd2.numpy(), idx1.numpy(),
                                 idx2.numpy())
        return torch.from_numpy(g1), torch.from_numpy(g2)


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, b, n, m, out):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from pointcloudregistration_b200 import sharded, synthetic

        x1, x2 = synthetic.make_pair("crosscopy", b, n, m, 60)
        if rank != 0:  # only the source holds real data; the others receive it by broadcast
            x1, x2 = torch.zeros_like(x1), torch.zeros_like(x2)
        a = x1.clone().requires_grad_(True)
        c = x2.clone().requires_grad_(True)
        with torch.no_grad():
            dist.broadcast(a, 0)
            dist.broadcast(c, 0)
        d1, d2, i1, i2 = sharded.nnd_sharded_with_idx(a, c, src=None, compute=OracleCompute)
        gd1, gd2 = synthetic.grad_pair(b, n, m, 60)
        ((d1 * gd1).sum() + (d2 * gd2).sum()).backward()
        # deterministic mode: replicated full backward from the gathered idx / graddist, no collective
        a2 = a.detach().clone().requires_grad_(True)
        c2 = c.detach().clone().requires_grad_(True)
        e1, e2, _, _ = sharded.nnd_sharded_with_idx(a2, c2, src=None, compute=OracleCompute, exact=True)
        ((e1 * gd1).sum() + (e2 * gd2).sum()).backward()
        # ... and from shards (gathered=False): graddist / idx are all-gathered first
        f1, f2, j1s, j2s, info_s = sharded.sharded_forward(a.detach(), c.detach(), gather=False, compute=OracleCompute)
        x_g1, x_g2 = sharded.sharded_backward(a.detach(), c.detach(), gd1[:, info_s.lo1:info_s.hi1], gd2[:, info_s.lo2:info_s.hi2],
                                              j1s, j2s, info_s, gathered=False, compute=OracleCompute, exact=True)
        torch.save(dict(d1=d1.detach(), d2=d2.detach(), i1=i1, i2=i2, g1=a.grad, g2=c.grad, e1=a2.grad, e2=c2.grad,
                        x1=x_g1, x2=x_g2), f"{out}/r{rank}.pt")
        # shard-only mode + explicit broadcast from src
        y1, y2 = synthetic.make_pair("crosscopy", b, n, m, 60)
        if rank != 0:
            y1, y2 = torch.zeros_like(y1), torch.zeros_like(y2)
        s1, s2, j1, j2, info = sharded.sharded_forward(y1, y2, src=0, gather=False, compute=OracleCompute)
        assert s1.shape == (b, info.hi1 - info.lo1) and s2.shape == (b, info.hi2 - info.lo2)
        assert torch.equal(s1, d1.detach()[:, info.lo1:info.hi1]) and torch.equal(j2, i2[:, info.lo2:info.hi2])
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world,b,n,m", [(2, 1, 301, 200), (2, 3, 64, 129), (3, 1, 100, 101), (3, 2, 10, 7)])
def test_sharded_matches_unsharded(world, b, n, m, tmp_path):
    import oracle
    from _util import assert_forward_exact, assert_grad_close
    from pointcloudregistration_b200 import synthetic

    mp.spawn(_worker, args=(world, _free_port(), b, n, m, str(tmp_path)), nprocs=world, join=True)
    x1, x2 = synthetic.make_pair("crosscopy", b, n, m, 60)
    gd1, gd2 = synthetic.grad_pair(b, n, m, 60)
    want = oracle.forward(x1.numpy(), x2.numpy())
    wg = oracle.backward(x1.numpy(), x2.numpy(), gd1.numpy(), gd2.numpy(), want[2], want[3])
    for r in range(world):
        z = torch.load(f"{tmp_path}/r{r}.pt")
        assert_forward_exact(tuple(z[k].numpy() for k in ("d1", "d2", "i1", "i2")), want, what=f"rank {r}: ")
        assert_grad_close((z["g1"].numpy(), z["g2"].numpy()), wg, what=f"rank {r}: ")
        for k1, k2 in (("e1", "e2"), ("x1", "x2")):  # deterministic mode: every bit, on every rank
            assert z[k1].numpy().tobytes() == wg[0].tobytes() and z[k2].numpy().tobytes() == wg[1].tobytes(), (r, k1)


def test_shard_ranges_cover_and_balance():
    from pointcloudregistration_b200.sharded import shard_range

    for n in (0, 1, 7, 8, 1_000_000, 1_000_003):
        for w in (1, 2, 3, 4, 8):
            for align in (1, 8):
                r = [shard_range(n, w, k, align) for k in range(w)]
                assert r[0][0] == 0 and r[-1][1] == n
                assert all(r[k][1] == r[k + 1][0] for k in range(w - 1))
                assert all(lo % align == 0 for lo, _ in r)
                sizes = [hi - lo for lo, hi in r]
                assert min(sizes) >= 0 and max(sizes) - min(sizes) <= (1 if align == 1 else 2 * align)


def test_uniform_shard_ranges_have_one_stride():
    """Cloud-1 row shards: a common stride, so the all_gather of the shards is the full array cut at n."""
    from pointcloudregistration_b200.sharded import shard_stride, uniform_shard_range

    for n in (0, 1, 7, 8, 10, 1_000_000, 1_000_003):
        for w in (1, 2, 3, 4, 8):
            for align in (1, 8):
                stride = shard_stride(n, w, align)
                r = [uniform_shard_range(n, w, k, align) for k in range(w)]
                assert stride % align == 0 and stride * w >= n
                assert r[0][0] == 0 and r[-1][1] == n and all(r[k][1] == r[k + 1][0] for k in range(w - 1))
                assert all(lo == min(n, k * stride) and hi - lo <= stride for k, (lo, hi) in enumerate(r))
    assert uniform_shard_range(1_000_000, 8, 3, 8) == (375_000, 500_000)


def test_single_process_without_process_group():
    from pointcloudregistration_b200 import sharded, synthetic
    import oracle

    x1, x2 = synthetic.make_pair("uniform", 2, 50, 60, 61)
    d1, d2, i1, i2, info = sharded.sharded_forward(x1, x2, compute=OracleCompute)
    assert info.world == 1
    want = oracle.forward(x1.numpy(), x2.numpy())
    assert np.array_equal(d1.numpy(), want[0]) and np.array_equal(i2.numpy(), want[3])
