"""Query-sharded nnd across the GPUs of one box (one process per GPU, torch.distributed / NCCL).

The reference has no multi-device path; this is new design for BASELINE.json
configs[3].  Every query row is independent, so rank r owns rows [r*N/G, (r+1)*N/G) of cloud 1
AND rows [r*M/G, (r+1)*M/G) of cloud 2 as *queries*; both clouds are needed in full on every
rank as *targets* (12 MB per million points), so they are broadcast over NVLink.  Exchange steps:

    forward : broadcast(xyz1), broadcast(xyz2)            [optional: caller may already hold them]
              all_gather(dist/idx shards)                 [optional: gather=False keeps shards]
    backward: all_reduce(SUM) of gradxyz1, gradxyz2       (scatter terms land anywhere in the other cloud)

Forward results are bit-identical for any G (pure per-row function).  Gradients differ from the
1-GPU result only by fp32 summation order (within the 1e-5 contract).

The collectives are plain NCCL calls: at N=M=1M they move ~60 MB per step against >= 50 ms of
compute per GPU, i.e. < 0.2 % of the step, so a fused compute+collective kernel has nothing to
overlap (numbers in DESIGN.md).

`compute` is the per-rank engine.  The product default is the CUDA C ABI; the CPU `gloo` tests
inject the oracle there to exercise this host logic without a GPU.
"""
from __future__ import annotations

from dataclasses import dataclass

import torch
import torch.distributed as dist
from torch.autograd import Function


def shard_range(n: int, world: int, rank: int) -> tuple[int, int]:
    """Contiguous balanced split of n rows: sizes differ by at most one."""
    return (n * rank) // world, (n * (rank + 1)) // world


@dataclass
class ShardInfo:
    rank: int
    world: int
    lo1: int
    hi1: int
    lo2: int
    hi2: int


class CudaCompute:
    """Per-rank engine on the C ABI (include/nnd_b200.h).  No fallback."""

    @staticmethod
    def query(q: torch.Tensor, t: torch.Tensor):
        from .functional import nnd_query_raw

        return nnd_query_raw(q, t)

    @staticmethod
    def backward_term(q, t, gd, idx, grad_q, grad_t) -> None:
        """grad_q[row] += 2*gd*(q - t[idx]); grad_t[idx] -= the same.  All tensors contiguous."""
        from . import _cabi
        from .functional import _stream_ptr

        lib = _cabi.load()
        b, nq, _ = q.shape
        nt = t.size(1)
        if b == 0 or nq == 0:
            return
        with torch.cuda.device(q.device):
            rc = lib.nnd_backward_term_cuda_f32(q.data_ptr(), t.data_ptr(), gd.data_ptr(), idx.data_ptr(), b, nq, nt,
                                                grad_q.data_ptr(), grad_t.data_ptr(), _stream_ptr(q.device))
        _cabi.check(rc, "nnd_backward_term_cuda_f32")


def _world(group):
    if not dist.is_available() or not dist.is_initialized():
        return 0, 1
    return dist.get_rank(group), dist.get_world_size(group)


def _gather_rows(shard: torch.Tensor, n: int, world: int, group) -> torch.Tensor:
    """all_gather of [B, rows_r] shards with rows_r = shard_range sizes (differ by <= 1) -> [B, n]."""
    if world == 1:
        return shard
    b = shard.size(0)
    longest = max(shard_range(n, world, r)[1] - shard_range(n, world, r)[0] for r in range(world))
    padded = shard.new_zeros(b, longest)
    padded[:, :shard.size(1)] = shard
    bufs = [torch.empty_like(padded) for _ in range(world)]
    dist.all_gather(bufs, padded, group=group)
    parts = []
    for r in range(world):
        lo, hi = shard_range(n, world, r)
        parts.append(bufs[r][:, :hi - lo])
    return torch.cat(parts, dim=1)


def sharded_forward(xyz1, xyz2, group=None, src=None, gather=True, compute=CudaCompute):
    """Returns (dist1, dist2, idx1, idx2, info).  With gather=True the four tensors are full-size and
    identical on every rank; otherwise they are this rank's row shards."""
    rank, world = _world(group)
    if src is not None and world > 1:
        dist.broadcast(xyz1, src=src, group=group)
        dist.broadcast(xyz2, src=src, group=group)
    n, m = xyz1.size(1), xyz2.size(1)
    lo1, hi1 = shard_range(n, world, rank)
    lo2, hi2 = shard_range(m, world, rank)
    info = ShardInfo(rank, world, lo1, hi1, lo2, hi2)
    d1, i1 = compute.query(xyz1[:, lo1:hi1].contiguous(), xyz2)
    d2, i2 = compute.query(xyz2[:, lo2:hi2].contiguous(), xyz1)
    if gather:
        d1, i1 = _gather_rows(d1, n, world, group), _gather_rows(i1, n, world, group)
        d2, i2 = _gather_rows(d2, m, world, group), _gather_rows(i2, m, world, group)
    return d1, d2, i1, i2, info


def sharded_backward(xyz1, xyz2, graddist1, graddist2, idx1, idx2, info: ShardInfo, group=None, gathered=True,
                     compute=CudaCompute):
    """Full-size gradxyz1, gradxyz2, identical on every rank (all_reduce SUM of per-rank partials).
    graddist*/idx* are full-size when gathered=True, else this rank's shards."""
    lo1, hi1, lo2, hi2 = info.lo1, info.hi1, info.lo2, info.hi2
    if gathered:
        graddist1, idx1 = graddist1[:, lo1:hi1], idx1[:, lo1:hi1]
        graddist2, idx2 = graddist2[:, lo2:hi2], idx2[:, lo2:hi2]
    g1 = torch.zeros_like(xyz1)
    g2 = torch.zeros_like(xyz2)
    own1 = torch.zeros(xyz1.size(0), hi1 - lo1, 3, dtype=xyz1.dtype, device=xyz1.device)
    own2 = torch.zeros(xyz2.size(0), hi2 - lo2, 3, dtype=xyz2.dtype, device=xyz2.device)
    compute.backward_term(xyz1[:, lo1:hi1].contiguous(), xyz2, graddist1.contiguous(), idx1.contiguous(), own1, g2)
    compute.backward_term(xyz2[:, lo2:hi2].contiguous(), xyz1, graddist2.contiguous(), idx2.contiguous(), own2, g1)
    g1[:, lo1:hi1] += own1
    g2[:, lo2:hi2] += own2
    if info.world > 1:
        dist.all_reduce(g1, op=dist.ReduceOp.SUM, group=group)
        dist.all_reduce(g2, op=dist.ReduceOp.SUM, group=group)
    return g1, g2


class ShardedNNDFunction(Function):
    """nnd with the query rows sharded over the ranks of `group`; inputs and outputs replicated."""

    @staticmethod
    def forward(ctx, xyz1, xyz2, group, src, compute):
        d1, d2, i1, i2, info = sharded_forward(xyz1, xyz2, group=group, src=src, gather=True, compute=compute)
        ctx.save_for_backward(xyz1, xyz2, i1, i2)
        ctx.info, ctx.group, ctx.compute = info, group, compute
        ctx.mark_non_differentiable(i1, i2)
        return d1, d2, i1, i2

    @staticmethod
    def backward(ctx, gd1, gd2, _a, _b):
        xyz1, xyz2, i1, i2 = ctx.saved_tensors
        g1, g2 = sharded_backward(xyz1, xyz2, gd1.contiguous(), gd2.contiguous(), i1, i2, ctx.info, group=ctx.group,
                                  gathered=True, compute=ctx.compute)
        return g1, g2, None, None, None


def nnd_sharded(xyz1, xyz2, group=None, src=None, compute=CudaCompute):
    """dist1, dist2 = nnd_sharded(xyz1, xyz2): same contract as nnd(), computed by all ranks together."""
    d1, d2, _, _ = ShardedNNDFunction.apply(xyz1, xyz2, group, src, compute)
    return d1, d2


def nnd_sharded_with_idx(xyz1, xyz2, group=None, src=None, compute=CudaCompute):
    return ShardedNNDFunction.apply(xyz1, xyz2, group, src, compute)
