"""Row-sharded nnd across the GPUs of one box (one process per GPU, torch.distributed / NCCL).

The reference has no multi-device path; this is new design for BASELINE.json
configs[3].  The forward evaluates every (xyz1[j], xyz2[k]) distance ONCE and serves both
directions from it (csrc/nnd_sym.cuh), so the unit of sharding is a range of xyz1 ROWS: rank r
owns rows [lo1_r, hi1_r) (boundaries on multiples of the kernel's row group).  Both clouds are
needed in full on every rank (12 MB per million points), so they are broadcast over NVLink.

    forward : broadcast(xyz1), broadcast(xyz2)            [optional: caller may already hold them]
              rows kernel  -> dist1/idx1 of the own rows (final) + per-column partial keys
              all_reduce(MIN, int64) of the column keys   <- the one real exchange step (8 B / point)
              column finalize -> dist2/idx2, replicated
              all_gather(dist1/idx1 row shards)           [optional: gather=False keeps shards]
    backward: term 1 over the own xyz1 rows, term 2 over the own slice of xyz2 rows,
              all_reduce(SUM) of gradxyz1, gradxyz2       (scatter terms land anywhere in the other cloud)
    backward, deterministic mode (functional.set_exact_backward / torch.use_deterministic_algorithms):
              every rank already holds the full idx / graddist, so each runs the whole deterministic backward
              (0.4 ms at 1M points) -- no collective, gradients bit-identical on every rank, for any G, and to
              the reference CPU path

Forward results are bit-identical for any G: row results are a pure per-row function, and a column
key (dist_bits << 32 | row_group) is an order-independent integer minimum.  Gradients differ from
the 1-GPU result only by fp32 summation order (within the 1e-5 contract).

On CUDA with torch's symmetric memory the exchange steps run over PEER MEMORY (class PeerExchange below): the merge of
the column keys, the column finalize and the gather of all results are ONE kernel per rank, the gradient sum another
(include/nnd_b200.h: nnd_exchange_cols_p2p_cuda_f32, nnd_sum_p2p_cuda_f32).  The NCCL collectives listed above are the
fallback (no symmetric memory, the CPU tests, NND_B200_PEER_EXCHANGE=0); both give the same bits (numbers in DESIGN.md 5).

`compute` is the per-rank engine.  The product default is the CUDA C ABI; the CPU `gloo` tests
inject the oracle there to exercise this host logic without a GPU.
"""
from __future__ import annotations

import os
from dataclasses import dataclass

import torch
import torch.distributed as dist
from torch.autograd import Function


#: set to a list to get a timeline of one step: _mark() then appends (phase name, CUDA event recorded on the current stream
#: once that phase's work has been enqueued / waited for); tools/sharded_timeline.py turns it into per-phase milliseconds
trace = None


def _mark(name: str, ref: torch.Tensor) -> None:
    if trace is not None and ref.is_cuda:
        ev = torch.cuda.Event(enable_timing=True)
        ev.record(torch.cuda.current_stream(ref.device))
        trace.append((name, ev))


def shard_range(n: int, world: int, rank: int, align: int = 1) -> tuple[int, int]:
    """Contiguous balanced split of n rows; interior boundaries are rounded down to multiples of `align`."""
    def cut(r: int) -> int:
        if r >= world:
            return n
        return (n * r) // world // align * align

    return cut(rank), cut(rank + 1)


def uniform_shard_range(n: int, world: int, rank: int, align: int = 1) -> tuple[int, int]:
    """Rows [rank * rpr, min(n, (rank + 1) * rpr)) with one stride rpr = ceil(n / world) rounded up to `align` for every
    rank: the row shards of cloud 1.  With a common stride the all_gather of the shards IS the full array (rank-major
    concatenation, cut at n): no per-rank offsets, no re-packing.  The last ranks may get fewer rows (or none, for tiny
    clouds); at 1M rows and 8 ranks every shard is exactly 125 000 rows."""
    rpr = shard_stride(n, world, align)
    lo = min(n, rank * rpr)
    return lo, min(n, lo + rpr)


def shard_stride(n: int, world: int, align: int = 1) -> int:
    return max(align, -(-(-(-n // world)) // align) * align)  # ceil(ceil(n / world) / align) * align, at least align


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

    launches = 0  # kernels launched through this engine so far (bench.py's gpu_launches at N > 1)

    @classmethod
    def _count(cls) -> None:
        from . import _cabi

        cls.launches += int(_cabi.load().nnd_last_launch_count())

    @staticmethod
    def row_group() -> int:
        from .functional import row_group

        return row_group()

    supports_peer_exchange = True  # rows() can write into caller-owned (peer-addressable) buffers

    @staticmethod
    def rows(xyz1: torch.Tensor, xyz2: torch.Tensor, lo: int, hi: int, out=None):
        """(dist1 [B,hi-lo], idx1 [B,hi-lo], colkeys int64) for rows [lo, hi) of xyz1 against all of xyz2."""
        from .functional import nnd_forward_rows_raw

        res = nnd_forward_rows_raw(xyz1, xyz2, lo, hi, out=out)
        CudaCompute._count()
        return res

    @staticmethod
    def finalize_cols(xyz1: torch.Tensor, xyz2: torch.Tensor, colkeys: torch.Tensor):
        from .functional import nnd_finalize_cols_raw

        out = nnd_finalize_cols_raw(xyz1, xyz2, colkeys)
        CudaCompute._count()
        return out

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
        CudaCompute._count()

    @staticmethod
    def exact_default() -> bool:
        from .functional import exact_backward_enabled

        return exact_backward_enabled()

    @staticmethod
    def backward_exact(xyz1, xyz2, graddist1, graddist2, idx1, idx2):
        """Both full gradients, bit-identical to my_lib.cpp:91-130 (nnd_backward_exact_cuda_f32)."""
        from .functional import nnd_backward_raw

        out = nnd_backward_raw(xyz1, xyz2, graddist1, graddist2, idx1, idx2, exact=True)
        CudaCompute._count()
        return out


#: False = always the NCCL collectives (A/B measurements; NND_B200_PEER_EXCHANGE=0 in the environment does the same)
peer_exchange = os.environ.get("NND_B200_PEER_EXCHANGE", "1") != "0"


class PeerExchange:
    """The exchange steps of one (group, device, B, N, M) problem over PEER MEMORY instead of NCCL collectives
    (include/nnd_b200.h: nnd_exchange_cols_p2p_cuda_f32, nnd_sum_p2p_cuda_f32; csrc/nnd_sym.cuh).

    Every rank keeps its column keys, dist1 / idx1, dist2 / idx2 and the flat gradient buffer in ONE symmetric-memory
    allocation (torch.distributed._symmetric_memory: the same virtual layout on every rank, every rank can address every
    other rank's copy over NVLink / NVSwitch).  forward: the rows kernel writes straight into it; then barrier, ONE kernel
    (merge of the column keys + column finalize of the own slice + stores of dist2 / idx2 and of the own rows of
    dist1 / idx1 into every rank), barrier.  backward: the gradient terms accumulate into the flat buffer; barrier, ONE
    kernel (two-shot sum in rank order), barrier.  The barriers are torch's symmetric-memory barrier on the current stream.
    Results: the same bits as the NCCL formulation (an integer minimum; the sum is in rank order, identical on every
    rank)."""

    _cache: dict = {}
    _unavailable = False

    @classmethod
    def get(cls, xyz1, xyz2, group, world, rank, compute):
        if (not peer_exchange or cls._unavailable or world < 2 or not xyz1.is_cuda
                or not getattr(compute, "supports_peer_exchange", False)):
            return None
        b, n, m = xyz1.size(0), xyz1.size(1), xyz2.size(1)
        key = (id(group) if group is not None else 0, xyz1.device.index, b, n, m, world)
        px = cls._cache.get(key)
        if px is None:
            why = None
            try:
                px = cls(xyz1.device, group, world, rank, b, n, m)
            except Exception as exc:  # noqa: BLE001 -- no symmetric memory on this box / backend: NCCL collectives it is
                why = f"{type(exc).__name__}: {str(exc)[:200]}"
            # every rank takes the same path: one rank without the buffers sends ALL ranks to the collectives (a rank
            # waiting at a peer-memory barrier for one that went to NCCL would never return).  First use only.
            ok = torch.tensor([0 if px is None else 1], dtype=torch.int32, device=xyz1.device)
            dist.all_reduce(ok, op=dist.ReduceOp.MIN, group=group)
            if int(ok.item()) == 0:
                import sys

                cls._unavailable = True
                print(f"[nnd_b200] peer-memory exchange unavailable ({why or 'another rank could not set it up'}); "
                      "using NCCL collectives", file=sys.stderr)
                return None
            cls._cache[key] = px
        return px

    def __init__(self, dev, group, world, rank, b, n, m):
        import ctypes

        import torch.distributed._symmetric_memory as symm_mem

        from . import _cabi

        lib = _cabi.load()
        if world > lib.nnd_peer_max_ranks():
            raise RuntimeError(f"at most {lib.nnd_peer_max_ranks()} ranks")
        self.lib, self.dev, self.world, self.rank, self.b, self.n, self.m = lib, dev, world, rank, b, n, m
        cc = int(lib.nnd_colkeys_count(b, m))
        parts = (("colkeys", cc * 8), ("dist1", b * n * 4), ("idx1", b * n * 4), ("dist2", b * m * 4), ("idx2", b * m * 4),
                 ("flat", 3 * b * (n + m) * 4))
        offs, total = {}, 0
        for name, nbytes in parts:
            offs[name] = total
            total += -(-nbytes // 256) * 256
        self.buf = symm_mem.empty(total, dtype=torch.uint8, device=dev)
        self.hdl = symm_mem.rendezvous(self.buf, group if group is not None else dist.group.WORLD)
        bases = [int(p_) for p_ in self.hdl.buffer_ptrs]
        self.ptrs = {name: (ctypes.c_void_p * world)(*[base + off for base in bases]) for name, off in offs.items()}

        def view(name, nbytes, dtype, shape):
            return self.buf[offs[name]:offs[name] + nbytes].view(dtype).view(shape)

        self.colkeys = view("colkeys", cc * 8, torch.int64, (cc,))
        self.dist1 = view("dist1", b * n * 4, torch.float32, (b, n))
        self.idx1 = view("idx1", b * n * 4, torch.int32, (b, n))
        self.dist2 = view("dist2", b * m * 4, torch.float32, (b, m))
        self.idx2 = view("idx2", b * m * 4, torch.int32, (b, m))
        self.flat = view("flat", 3 * b * (n + m) * 4, torch.float32, (3 * b * (n + m),))

    def exchange_cols(self, xyz1, xyz2, lo1: int, hi1: int) -> None:
        """After rows() wrote this rank's keys and rows into the buffers: every rank ends up with all of dist1, idx1,
        dist2, idx2."""
        from . import _cabi
        from .functional import _stream_ptr

        self.hdl.barrier(channel=0)  # every rank's rows kernel is done: its keys are complete
        with torch.cuda.device(self.dev):
            rc = self.lib.nnd_exchange_cols_p2p_cuda_f32(self.ptrs["colkeys"], self.ptrs["dist1"], self.ptrs["idx1"],
                                                         self.ptrs["dist2"], self.ptrs["idx2"], self.world, self.rank,
                                                         xyz1.data_ptr(), xyz2.data_ptr(), self.b, self.n, self.m,
                                                         int(lo1), int(hi1), _stream_ptr(self.dev))
        _cabi.check(rc, "nnd_exchange_cols_p2p_cuda_f32")
        CudaCompute._count()
        self.hdl.barrier(channel=0)  # every rank's stores have landed (and nobody rewrites its keys before all have read)

    def sum_flat(self) -> None:
        """In-place sum of every rank's flat gradient buffer (rank order: the same bits everywhere)."""
        from . import _cabi
        from .functional import _stream_ptr

        self.hdl.barrier(channel=1)
        with torch.cuda.device(self.dev):
            rc = self.lib.nnd_sum_p2p_cuda_f32(self.ptrs["flat"], self.world, self.rank, self.flat.numel(),
                                               _stream_ptr(self.dev))
        _cabi.check(rc, "nnd_sum_p2p_cuda_f32")
        CudaCompute._count()
        self.hdl.barrier(channel=1)


def _world(group):
    if not dist.is_available() or not dist.is_initialized():
        return 0, 1
    return dist.get_rank(group), dist.get_world_size(group)


def _gather_rows(shard: torch.Tensor, n: int, world: int, group, align: int) -> torch.Tensor:
    """all_gather of [B, rows_r] shards (rows_r = shard_range sizes) -> [B, n].  General form (any split): used for the
    cloud-2 slices of the deterministic backward's inputs."""
    if world == 1:
        return shard.contiguous()
    b = shard.size(0)
    ranges = [shard_range(n, world, r, align) for r in range(world)]
    longest = max(hi - lo for lo, hi in ranges)
    padded = shard.new_zeros(b, longest)
    padded[:, :shard.size(1)] = shard
    bufs = [torch.empty_like(padded) for _ in range(world)]
    dist.all_gather(bufs, padded, group=group)
    return torch.cat([bufs[r][:, :hi - lo] for r, (lo, hi) in enumerate(ranges)], dim=1)


def _gather_pair_start(a: torch.Tensor, c: torch.Tensor, n: int, world: int, group, align: int):
    """Starts ONE all_gather for two [B, rows] 4-byte shards of cloud-1 rows (uniform_shard_range: common stride).
    Send layout [2, B, stride], receive layout [world, 2, B, stride]; returns (work handle, receive buffer).
    Two small copies in, nothing zero-filled (the padding is never read)."""
    b, rows = a.shape
    stride = shard_stride(n, world, align)
    send = torch.empty(2, b, stride, dtype=torch.int32, device=a.device)
    send[0, :, :rows] = a.view(torch.int32)
    send[1, :, :rows] = c.view(torch.int32)
    recv = torch.empty(world * 2, b, stride, dtype=torch.int32, device=a.device)  # concatenation along dim 0 (the
    work = dist.all_gather_into_tensor(recv, send, group=group, async_op=True)    # form every backend accepts)
    return work, recv.view(world, 2, b, stride)


def _gather_pair_finish(work, recv: torch.Tensor, n: int, dtype_a, dtype_c):
    """Waits for the gather and returns the two full [B, n] tensors (own storage each: one strided copy per tensor)."""
    work.wait()
    world, _, b, stride = recv.shape
    out = []
    for which, dtype in ((0, dtype_a), (1, dtype_c)):
        # [world, B, stride] -> [B, world * stride]: rank-major rows; the chunks of one tensor are 2*B*stride apart in
        # recv, so this reshape is always a copy into storage of its own
        full = recv[:, which].permute(1, 0, 2).reshape(b, world * stride)
        out.append(full[:, :n].contiguous().view(dtype))  # no-op for one cloud pair (a [1, n] prefix is contiguous)
    return out[0], out[1]


def sharded_forward(xyz1, xyz2, group=None, src=None, gather=True, compute=CudaCompute):
    """Returns (dist1, dist2, idx1, idx2, info).  With gather=True the four tensors are full-size and
    identical on every rank; otherwise they are this rank's row shards ([lo1,hi1) of cloud 1, [lo2,hi2) of
    cloud 2)."""
    rank, world = _world(group)
    _mark("start", xyz1)
    if src is not None and world > 1:
        dist.broadcast(xyz1, src=src, group=group)
        dist.broadcast(xyz2, src=src, group=group)
        _mark("broadcast xyz1, xyz2", xyz1)
    n, m = xyz1.size(1), xyz2.size(1)
    align = compute.row_group()
    lo1, hi1 = uniform_shard_range(n, world, rank, align)
    lo2, hi2 = shard_range(m, world, rank)
    info = ShardInfo(rank, world, lo1, hi1, lo2, hi2)
    if xyz1.size(0) == 0 or n == 0 or m == 0:  # my_lib.cpp:9-10,21-22: empty inner loop -> zeros
        z = lambda cols: (xyz1.new_zeros(xyz1.size(0), cols), torch.zeros(xyz1.size(0), cols, dtype=torch.int32,  # noqa: E731
                                                                          device=xyz1.device))
        (d1, i1), (d2, i2) = z(n if gather else hi1 - lo1), z(m if gather else hi2 - lo2)
        return d1, d2, i1, i2, info
    px = PeerExchange.get(xyz1, xyz2, group, world, rank, compute) if gather else None
    if px is not None:  # the exchange over peer memory: one kernel instead of all_reduce + finalize + all_gather
        compute.rows(xyz1.contiguous(), xyz2.contiguous(), lo1, hi1, out=(px.dist1, px.idx1, px.colkeys))
        _mark("rows kernel (prepare + scan + finalize rows)", xyz1)
        px.exchange_cols(xyz1.contiguous(), xyz2.contiguous(), lo1, hi1)
        _mark("peer-memory exchange: barrier, merge + column finalize + stores to every rank (one kernel), barrier", xyz1)
        out = px.dist1.clone(), px.dist2.clone(), px.idx1.clone(), px.idx2.clone()  # the buffers are reused by the next call
        _mark("copies of the four results out of the exchange buffers", xyz1)
        return out[0], out[1], out[2], out[3], info
    d1, i1, colkeys = compute.rows(xyz1, xyz2, lo1, hi1)
    _mark("rows kernel (prepare + scan + finalize rows)", xyz1)
    pending = None
    if world > 1:
        # the one real exchange step first (the column finalize waits for it) ...
        kwork = dist.all_reduce(colkeys, op=dist.ReduceOp.MIN, group=group, async_op=True)
        if gather:  # ... then the row results: one collective, in flight while the column finalize runs
            pending = _gather_pair_start(d1.contiguous(), i1.contiguous(), n, world, group, align)
        kwork.wait()
        _mark("all_reduce(MIN) column keys", xyz1)
    d2, i2 = compute.finalize_cols(xyz1, xyz2, colkeys)
    _mark("finalize columns", xyz1)
    if gather:
        if pending is not None:
            d1, i1 = _gather_pair_finish(*pending, n, torch.float32, torch.int32)
            _mark("all_gather dist1/idx1 (issued before the column finalize) + unpack", xyz1)
        else:
            d1, i1 = d1.contiguous(), i1.contiguous()
    else:
        d1, i1 = d1.contiguous(), i1.contiguous()
        d2, i2 = d2[:, lo2:hi2].contiguous(), i2[:, lo2:hi2].contiguous()
    return d1, d2, i1, i2, info


def sharded_backward(xyz1, xyz2, graddist1, graddist2, idx1, idx2, info: ShardInfo, group=None, gathered=True,
                     compute=CudaCompute, exact=None):
    """Full-size gradxyz1, gradxyz2, identical on every rank (all_reduce SUM of per-rank partials).
    graddist*/idx* are full-size when gathered=True, else this rank's shards.
    exact (None = the engine's default, i.e. functional.exact_backward_enabled()): the deterministic backward,
    replicated on every rank from the full idx / graddist -- no collective, bit-identical gradients for any G."""
    lo1, hi1, lo2, hi2 = info.lo1, info.hi1, info.lo2, info.hi2
    if exact is None:
        exact = bool(getattr(compute, "exact_default", lambda: False)())
    if exact:
        if not gathered and info.world > 1:  # shards in: one all_gather per cloud (bits side by side, like the forward)
            n, m, b = xyz1.size(1), xyz2.size(1), xyz1.size(0)
            pending = _gather_pair_start(graddist1.contiguous(), idx1.contiguous(), n, info.world, group,
                                         compute.row_group())
            both2 = _gather_rows(torch.cat([graddist2.contiguous().view(torch.int32), idx2.contiguous()], dim=0), m,
                                 info.world, group, 1)
            graddist1, idx1 = _gather_pair_finish(*pending, n, torch.float32, torch.int32)
            graddist2, idx2 = both2[:b].clone().view(torch.float32), both2[b:].clone()
        return compute.backward_exact(xyz1, xyz2, graddist1.contiguous(), graddist2.contiguous(), idx1.contiguous(),
                                      idx2.contiguous())
    if gathered:
        graddist1, idx1 = graddist1[:, lo1:hi1], idx1[:, lo1:hi1]
        graddist2, idx2 = graddist2[:, lo2:hi2], idx2[:, lo2:hi2]
    # one flat buffer for both gradients: a single all_reduce (or one peer-memory sum) instead of two
    px = PeerExchange.get(xyz1, xyz2, group, info.world, info.rank, compute) if info.world > 1 else None
    if px is not None:
        flat = px.flat
        flat.zero_()
    else:
        flat = torch.zeros(xyz1.numel() + xyz2.numel(), dtype=xyz1.dtype, device=xyz1.device)
    g1 = flat[:xyz1.numel()].view_as(xyz1)
    g2 = flat[xyz1.numel():].view_as(xyz2)
    if xyz1.size(0) == 1:
        # one cloud pair (the large-cloud case): a row range of a [1, n, 3] tensor is contiguous, so the own-row terms
        # accumulate straight into the flat buffer -- no shard buffers, no copies, no extra adds
        compute.backward_term(xyz1[:, lo1:hi1], xyz2, graddist1.contiguous(), idx1.contiguous(), g1[:, lo1:hi1], g2)
        compute.backward_term(xyz2[:, lo2:hi2], xyz1, graddist2.contiguous(), idx2.contiguous(), g2[:, lo2:hi2], g1)
    else:
        own1 = torch.zeros(xyz1.size(0), hi1 - lo1, 3, dtype=xyz1.dtype, device=xyz1.device)
        own2 = torch.zeros(xyz2.size(0), hi2 - lo2, 3, dtype=xyz2.dtype, device=xyz2.device)
        compute.backward_term(xyz1[:, lo1:hi1].contiguous(), xyz2, graddist1.contiguous(), idx1.contiguous(), own1, g2)
        compute.backward_term(xyz2[:, lo2:hi2].contiguous(), xyz1, graddist2.contiguous(), idx2.contiguous(), own2, g1)
        g1[:, lo1:hi1] += own1
        g2[:, lo2:hi2] += own2
    _mark("backward: zero-fill + both gradient terms", xyz1)
    if px is not None:
        px.sum_flat()
        _mark("peer-memory sum of the gradients: barrier, two-shot sum (one kernel), barrier", xyz1)
        return g1.clone(), g2.clone()  # the exchange buffer is reused by the next call
    if info.world > 1:
        dist.all_reduce(flat, op=dist.ReduceOp.SUM, group=group)
        _mark("all_reduce(SUM) gradients", xyz1)
    return g1, g2


class ShardedNNDFunction(Function):
    """nnd with the query rows sharded over the ranks of `group`; inputs and outputs replicated."""

    @staticmethod
    def forward(ctx, xyz1, xyz2, group, src, compute, exact=None):
        d1, d2, i1, i2, info = sharded_forward(xyz1, xyz2, group=group, src=src, gather=True, compute=compute)
        ctx.save_for_backward(xyz1, xyz2, i1, i2)
        ctx.info, ctx.group, ctx.compute, ctx.exact = info, group, compute, exact
        ctx.mark_non_differentiable(i1, i2)
        return d1, d2, i1, i2

    @staticmethod
    def backward(ctx, gd1, gd2, _a, _b):
        xyz1, xyz2, i1, i2 = ctx.saved_tensors
        g1, g2 = sharded_backward(xyz1, xyz2, gd1.contiguous(), gd2.contiguous(), i1, i2, ctx.info, group=ctx.group,
                                  gathered=True, compute=ctx.compute, exact=ctx.exact)
        return g1, g2, None, None, None, None


def nnd_sharded(xyz1, xyz2, group=None, src=None, compute=CudaCompute, exact=None):
    """dist1, dist2 = nnd_sharded(xyz1, xyz2): same contract as nnd(), computed by all ranks together.
    exact: deterministic backward (None = follow functional.set_exact_backward / torch's deterministic mode)."""
    d1, d2, _, _ = ShardedNNDFunction.apply(xyz1, xyz2, group, src, compute, exact)
    return d1, d2


def nnd_sharded_with_idx(xyz1, xyz2, group=None, src=None, compute=CudaCompute, exact=None):
    return ShardedNNDFunction.apply(xyz1, xyz2, group, src, compute, exact)
