"""Host-side mirror of the reference's autograd boundary for nnd.

Mirrors dip/torch-nndistance/torch_nndistance/__init__.py:10-61 (NNDFunction,
nnd) on top of the C ABI in include/nnd_b200.h.  Same names, same argument
meaning, same 2-tuple return.  Differences, all deliberate:
outputs are allocated directly on xyz1's device (the reference builds them on
the CPU and copies, __init__.py:17-28), kernels run on torch's current stream
(the reference uses the legacy default stream, nnd_cuda.cu:152), inputs are
validated, and CPU tensors are rejected -- this build has no CPU path.
"""
from __future__ import annotations

import torch
from torch.autograd import Function

from . import _cabi


def _check_clouds(xyz1: torch.Tensor, xyz2: torch.Tensor) -> None:
    for name, t in (("xyz1", xyz1), ("xyz2", xyz2)):
        if not isinstance(t, torch.Tensor):
            raise TypeError(f"{name} must be a torch.Tensor")
        if t.dim() != 3 or t.size(2) != 3:
            raise ValueError(f"{name} must have shape [B, N, 3], got {tuple(t.shape)}")
        if t.dtype != torch.float32:
            raise TypeError(f"{name} must be float32 (my_lib.cpp:50-51 reads float), got {t.dtype}")
        if not t.is_cuda:
            raise RuntimeError(
                f"{name} is on {t.device}: pointcloudregistration_b200 is the B200 build of nnd and has no CPU path")
    if xyz1.size(0) != xyz2.size(0):
        raise ValueError(f"batch sizes differ: {xyz1.size(0)} vs {xyz2.size(0)}")
    if xyz1.device != xyz2.device:
        raise RuntimeError(f"xyz1 is on {xyz1.device} but xyz2 is on {xyz2.device}")


def _stream_ptr(device: torch.device) -> int:
    # the raw cudaStream_t of torch's current stream; ~10x cheaper than building a torch.cuda.Stream object
    return torch._C._cuda_getCurrentRawStream(device.index if device.index is not None else torch.cuda.current_device())


def _workspace(nbytes: int, device: torch.device) -> torch.Tensor:
    # torch's caching allocator hands out >=512-byte aligned blocks
    return torch.empty(max(int(nbytes), 256), dtype=torch.uint8, device=device)


class _on_device:
    """`with torch.cuda.device(dev)` only when dev is not already current: the guard costs several microseconds per
    call, which is most of a small-cloud forward (DIP patches, NDP clouds)."""

    __slots__ = ("guard",)

    def __init__(self, dev: torch.device):
        idx = dev.index
        self.guard = None if idx is None or idx == torch.cuda.current_device() else torch.cuda.device(dev)

    def __enter__(self):
        if self.guard is not None:
            self.guard.__enter__()

    def __exit__(self, *exc):
        if self.guard is not None:
            self.guard.__exit__(*exc)


_search_mode = "brute"  # the reference's algorithm and this project's metric; "pruned" is opt-in
#: below this many points per cloud / this many distance evaluations per call the pruned search's fixed cost (two
#: sorts, tile boxes: ~0.3 ms) is not worth it -- measured crossover, profiles/r02_pruned_time_v3.txt: a lone 16 384-point
#: pair 0.16 ms brute-force against 0.32 ms pruned, C2 (32 x 16 384) 2.68 against 0.95 ms, C3 (100 k) 2.87 against 0.67 ms
PRUNED_MIN_POINTS = 2048
PRUNED_MIN_EVALS = 2_000_000_000


def set_search_mode(mode: str) -> str:
    """'brute' (default): every forward evaluates all pairs (the reference's algorithm, my_lib.cpp:3-25).
    'pruned': forwards of at least PRUNED_MIN_EVALS distance evaluations (clouds of at least PRUNED_MIN_POINTS points) take
    nnd_forward_pruned_cuda_f32 --
    the same bits (dist, idx, ties, NaN rules), one to two orders of magnitude faster on overlapping clouds.
    Returns the previous mode."""
    global _search_mode
    if mode not in ("brute", "pruned"):
        raise ValueError("search mode must be 'brute' or 'pruned'")
    before, _search_mode = _search_mode, mode
    return before


def search_mode() -> str:
    return _search_mode


def _pruned_route(b: int, n: int, m: int) -> bool:
    return (_search_mode == "pruned" and min(n, m) >= PRUNED_MIN_POINTS and b <= 65535
            and b * n * m >= PRUNED_MIN_EVALS and b * max(n, m) < 2 ** 31)


def nnd_forward_pruned_raw(xyz1: torch.Tensor, xyz2: torch.Tensor):
    """dist1, dist2, idx1, idx2 like nnd_forward_raw, bit-identical, through the exact pruned search
    (include/nnd_b200.h: nnd_forward_pruned_cuda_f32; csrc/nnd_pruned.cuh)."""
    _check_clouds(xyz1, xyz2)
    lib = _cabi.load()
    xyz1 = xyz1.contiguous()
    xyz2 = xyz2.contiguous()
    b, n, _ = xyz1.shape
    m = xyz2.size(1)
    dev = xyz1.device
    dist1 = torch.empty(b, n, dtype=torch.float32, device=dev)
    dist2 = torch.empty(b, m, dtype=torch.float32, device=dev)
    idx1 = torch.empty(b, n, dtype=torch.int32, device=dev)
    idx2 = torch.empty(b, m, dtype=torch.int32, device=dev)
    with _on_device(dev):
        nbytes = lib.nnd_pruned_workspace_bytes(b, n, m)
        if nbytes == 0:
            raise ValueError("nnd_forward_pruned: b*n and b*m must stay below 2^31")
        ws = _workspace(nbytes, dev)
        rc = lib.nnd_forward_pruned_cuda_f32(xyz1.data_ptr(), xyz2.data_ptr(), b, n, m, dist1.data_ptr(), dist2.data_ptr(),
                                             idx1.data_ptr(), idx2.data_ptr(), ws.data_ptr(), ws.numel(), _stream_ptr(dev))
    if rc:
        _cabi.check(rc, "nnd_forward_pruned_cuda_f32")
    return dist1, dist2, idx1, idx2


def nnd_forward_raw(xyz1: torch.Tensor, xyz2: torch.Tensor, hint_idx2: torch.Tensor = None):
    """dist1[B,N], dist2[B,M] float32 and idx1[B,N], idx2[B,M] int32; no autograd.
    hint_idx2 (optional, int32 [B,M]): the idx2 of a previous call on nearly the same clouds (a registration loop's last
    iteration) -- a warm start for the large-cloud kernels (include/nnd_b200.h: nnd_forward_warm_cuda_f32).  The
    results never depend on it."""
    _check_clouds(xyz1, xyz2)
    if _pruned_route(xyz1.size(0), xyz1.size(1), xyz2.size(1)):
        return nnd_forward_pruned_raw(xyz1, xyz2)
    lib = _cabi.load()
    xyz1 = xyz1.contiguous()
    xyz2 = xyz2.contiguous()
    b, n, _ = xyz1.shape
    m = xyz2.size(1)
    dev = xyz1.device
    dist1 = torch.empty(b, n, dtype=torch.float32, device=dev)
    dist2 = torch.empty(b, m, dtype=torch.float32, device=dev)
    idx1 = torch.empty(b, n, dtype=torch.int32, device=dev)
    idx2 = torch.empty(b, m, dtype=torch.int32, device=dev)
    with _on_device(dev):
        nbytes = lib.nnd_workspace_bytes(b, n, m)
        if nbytes == 0:  # small clouds: one launch, no workspace (include/nnd_b200.h)
            rc = lib.nnd_forward_cuda_f32(xyz1.data_ptr(), xyz2.data_ptr(), b, n, m, dist1.data_ptr(),
                                          dist2.data_ptr(), idx1.data_ptr(), idx2.data_ptr(), None, 0, _stream_ptr(dev))
        else:
            ws = _workspace(nbytes, dev)
            hint = None
            if hint_idx2 is not None:
                if hint_idx2.dtype != torch.int32 or hint_idx2.shape != (b, m) or hint_idx2.device != dev:
                    raise ValueError("hint_idx2 must be an int32 [B, M] tensor on the clouds' device")
                hint = hint_idx2.contiguous().data_ptr()
            rc = lib.nnd_forward_warm_cuda_f32(xyz1.data_ptr(), xyz2.data_ptr(), b, n, m, dist1.data_ptr(),
                                               dist2.data_ptr(), idx1.data_ptr(), idx2.data_ptr(), hint, ws.data_ptr(),
                                               ws.numel(), _stream_ptr(dev))
    if rc:
        _cabi.check(rc, "nnd_forward_cuda_f32")
    return dist1, dist2, idx1, idx2


def _len_ptr(lengths, b: int, dev):
    """int64 device tensor of per-cloud lengths (kept alive by the caller) or None."""
    if lengths is None:
        return None
    if lengths.dtype != torch.int64 or lengths.device != dev or lengths.shape != (b,):
        lengths = lengths.to(device=dev, dtype=torch.int64).reshape(b)
    return lengths.contiguous()


def ragged_supported(n: int, m: int) -> bool:
    """True when clouds of n and m points take the single-launch kernels, which handle per-cloud lengths themselves."""
    return bool(_cabi.load().nnd_ragged_supported(int(n), int(m)))


def nnd_forward_ragged_raw(xyz1, xyz2, len1=None, len2=None):
    """Both directions with per-cloud valid lengths (int64 tensors [B] or None) in ONE launch for the whole batch
    (include/nnd_b200.h: nnd_forward_ragged_cuda_f32).  Rows at or past a length come back as dist 0 / idx 0."""
    _check_clouds(xyz1, xyz2)
    lib = _cabi.load()
    xyz1, xyz2 = xyz1.contiguous(), xyz2.contiguous()
    b, n, _ = xyz1.shape
    m = xyz2.size(1)
    dev = xyz1.device
    len1, len2 = _len_ptr(len1, b, dev), _len_ptr(len2, b, dev)
    dist1 = torch.empty(b, n, dtype=torch.float32, device=dev)
    dist2 = torch.empty(b, m, dtype=torch.float32, device=dev)
    idx1 = torch.empty(b, n, dtype=torch.int32, device=dev)
    idx2 = torch.empty(b, m, dtype=torch.int32, device=dev)
    with _on_device(dev):
        rc = lib.nnd_forward_ragged_cuda_f32(xyz1.data_ptr(), xyz2.data_ptr(), b, n, m,
                                             None if len1 is None else len1.data_ptr(),
                                             None if len2 is None else len2.data_ptr(), dist1.data_ptr(),
                                             dist2.data_ptr(), idx1.data_ptr(), idx2.data_ptr(), _stream_ptr(dev))
    _cabi.check(rc, "nnd_forward_ragged_cuda_f32")
    return dist1, dist2, idx1, idx2


def nnd_query_ragged_raw(query, target, lenq=None, lent=None):
    """One direction with per-cloud lengths, one launch (nnd_query_ragged_cuda_f32)."""
    _check_clouds(query, target)
    lib = _cabi.load()
    query, target = query.contiguous(), target.contiguous()
    b, nq, _ = query.shape
    nt = target.size(1)
    dev = query.device
    lenq, lent = _len_ptr(lenq, b, dev), _len_ptr(lent, b, dev)
    dist = torch.empty(b, nq, dtype=torch.float32, device=dev)
    idx = torch.empty(b, nq, dtype=torch.int32, device=dev)
    with _on_device(dev):
        rc = lib.nnd_query_ragged_cuda_f32(query.data_ptr(), target.data_ptr(), b, nq, nt,
                                           None if lenq is None else lenq.data_ptr(),
                                           None if lent is None else lent.data_ptr(), dist.data_ptr(), idx.data_ptr(),
                                           _stream_ptr(dev))
    _cabi.check(rc, "nnd_query_ragged_cuda_f32")
    return dist, idx


def nnd_backward_ragged_raw(xyz1, xyz2, graddist1, graddist2, idx1, idx2, len1=None, len2=None):
    """gradxyz1, gradxyz2 of the ragged forward, one launch (nnd_backward_ragged_cuda_f32); rows past a length get 0."""
    lib = _cabi.load()
    b, n, _ = xyz1.shape
    m = xyz2.size(1)
    dev = xyz1.device
    len1, len2 = _len_ptr(len1, b, dev), _len_ptr(len2, b, dev)
    graddist1 = graddist1.to(device=dev, dtype=torch.float32).contiguous()
    graddist2 = graddist2.to(device=dev, dtype=torch.float32).contiguous()
    g1, g2 = torch.empty_like(xyz1), torch.empty_like(xyz2)
    with _on_device(dev):
        rc = lib.nnd_backward_ragged_cuda_f32(xyz1.data_ptr(), xyz2.data_ptr(), graddist1.data_ptr(),
                                              graddist2.data_ptr(), idx1.data_ptr(), idx2.data_ptr(), b, n, m,
                                              None if len1 is None else len1.data_ptr(),
                                              None if len2 is None else len2.data_ptr(), g1.data_ptr(), g2.data_ptr(),
                                              _stream_ptr(dev))
    _cabi.check(rc, "nnd_backward_ragged_cuda_f32")
    return g1, g2


def nnd_knn_backward_ragged_raw(query, target, graddist, idx, lenq=None, lent=None):
    """grad_query, grad_target of the one-direction ragged query, one launch (nnd_knn_backward_ragged_cuda_f32)."""
    lib = _cabi.load()
    b, nq, _ = query.shape
    nt = target.size(1)
    dev = query.device
    lenq, lent = _len_ptr(lenq, b, dev), _len_ptr(lent, b, dev)
    graddist = graddist.to(device=dev, dtype=torch.float32).contiguous()
    gq, gt = torch.empty_like(query), torch.empty_like(target)
    with _on_device(dev):
        rc = lib.nnd_knn_backward_ragged_cuda_f32(query.data_ptr(), target.data_ptr(), graddist.data_ptr(),
                                                  idx.data_ptr(), b, nq, nt, None if lenq is None else lenq.data_ptr(),
                                                  None if lent is None else lent.data_ptr(), gq.data_ptr(),
                                                  gt.data_ptr(), _stream_ptr(dev))
    _cabi.check(rc, "nnd_knn_backward_ragged_cuda_f32")
    return gq, gt


def row_group() -> int:
    """Row ranges handed to nnd_forward_rows_raw must start (and, unless they end at N, end) on a multiple of this."""
    return int(_cabi.load().nnd_row_group())


def nnd_forward_rows_raw(xyz1: torch.Tensor, xyz2: torch.Tensor, row_lo: int, row_hi: int, out=None):
    """Symmetric forward over rows [row_lo, row_hi) of xyz1 (include/nnd_b200.h: nnd_forward_rows_cuda_f32).

    Returns (dist1 [B, rows] f32, idx1 [B, rows] int32, colkeys [B * nnd_colkeys_count(1, M)] int64): the rows'
    final results and the packed per-column partial results, to be merged across ranks with an integer MIN and
    resolved by nnd_finalize_cols_raw.
    out (optional): (dist1 [B, N] f32, idx1 [B, N] int32, colkeys int64) caller-owned FULL-SIZE buffers to write into
    (the multi-GPU path passes peer-addressable memory, sharded.PeerExchange)."""
    _check_clouds(xyz1, xyz2)
    lib = _cabi.load()
    xyz1 = xyz1.contiguous()
    xyz2 = xyz2.contiguous()
    b, n, _ = xyz1.shape
    m = xyz2.size(1)
    dev = xyz1.device
    if b == 0 or n == 0 or m == 0:
        raise ValueError("nnd_forward_rows_raw needs non-empty clouds")
    if out is not None:
        dist1, idx1, colkeys = out
        if (dist1.shape != (b, n) or idx1.shape != (b, n) or dist1.dtype != torch.float32 or idx1.dtype != torch.int32
                or colkeys.dtype != torch.int64 or colkeys.numel() != lib.nnd_colkeys_count(b, m)
                or not (dist1.is_contiguous() and idx1.is_contiguous() and colkeys.is_contiguous())
                or dist1.device != dev or idx1.device != dev or colkeys.device != dev):
            raise ValueError("nnd_forward_rows_raw: `out` buffers have the wrong shape, dtype or device")
    else:
        dist1 = torch.empty(b, n, dtype=torch.float32, device=dev)
        idx1 = torch.empty(b, n, dtype=torch.int32, device=dev)
    with torch.cuda.device(dev):
        if out is None:
            colkeys = torch.empty(lib.nnd_colkeys_count(b, m), dtype=torch.int64, device=dev)
        ws = _workspace(lib.nnd_rows_workspace_bytes(b, n, m), dev)
        rc = lib.nnd_forward_rows_cuda_f32(xyz1.data_ptr(), xyz2.data_ptr(), b, n, m, int(row_lo), int(row_hi),
                                           dist1.data_ptr(), idx1.data_ptr(), colkeys.data_ptr(), ws.data_ptr(),
                                           ws.numel(), _stream_ptr(dev))
    _cabi.check(rc, "nnd_forward_rows_cuda_f32")
    return dist1[:, row_lo:row_hi], idx1[:, row_lo:row_hi], colkeys


def nnd_finalize_cols_raw(xyz1: torch.Tensor, xyz2: torch.Tensor, colkeys: torch.Tensor):
    """dist2 [B,M] f32, idx2 [B,M] int32 from (merged) column keys."""
    _check_clouds(xyz1, xyz2)
    lib = _cabi.load()
    xyz1 = xyz1.contiguous()
    xyz2 = xyz2.contiguous()
    b, n, _ = xyz1.shape
    m = xyz2.size(1)
    dev = xyz1.device
    if colkeys.dtype != torch.int64 or not colkeys.is_contiguous() or colkeys.device != dev:
        raise TypeError("colkeys must be the contiguous int64 tensor nnd_forward_rows_raw returned")
    if colkeys.numel() != lib.nnd_colkeys_count(b, m):
        raise ValueError("colkeys has the wrong length for these clouds")
    dist2 = torch.empty(b, m, dtype=torch.float32, device=dev)
    idx2 = torch.empty(b, m, dtype=torch.int32, device=dev)
    with torch.cuda.device(dev):
        rc = lib.nnd_finalize_cols_cuda_f32(xyz1.data_ptr(), xyz2.data_ptr(), b, n, m, colkeys.data_ptr(),
                                            dist2.data_ptr(), idx2.data_ptr(), _stream_ptr(dev))
    _cabi.check(rc, "nnd_finalize_cols_cuda_f32")
    return dist2, idx2


def nnd_query_raw(query: torch.Tensor, target: torch.Tensor):
    """One direction: for every query point its nearest target. dist[B,Nq] f32, idx[B,Nq] int32."""
    _check_clouds(query, target)
    lib = _cabi.load()
    query = query.contiguous()
    target = target.contiguous()
    b, nq, _ = query.shape
    nt = target.size(1)
    dev = query.device
    dist = torch.empty(b, nq, dtype=torch.float32, device=dev)
    idx = torch.empty(b, nq, dtype=torch.int32, device=dev)
    if _pruned_route(b, nq, nt) and 0 < (pbytes := lib.nnd_pruned_workspace_bytes(b, nq, nt)):
        with _on_device(dev):  # the exact pruned search, one direction (set_search_mode("pruned")): same bits
            ws = _workspace(pbytes, dev)
            rc = lib.nnd_query_pruned_cuda_f32(query.data_ptr(), target.data_ptr(), b, nq, nt, dist.data_ptr(),
                                               idx.data_ptr(), ws.data_ptr(), ws.numel(), _stream_ptr(dev))
        _cabi.check(rc, "nnd_query_pruned_cuda_f32")
        return dist, idx
    with _on_device(dev):
        nbytes = lib.nnd_query_workspace_bytes(b, nq, nt)
        ws = _workspace(nbytes, dev) if nbytes else None
        rc = lib.nnd_query_cuda_f32(query.data_ptr(), target.data_ptr(), b, nq, nt, dist.data_ptr(), idx.data_ptr(),
                                    ws.data_ptr() if nbytes else None, ws.numel() if nbytes else 0, _stream_ptr(dev))
    _cabi.check(rc, "nnd_query_cuda_f32")
    return dist, idx


_exact_backward = None  # None: follow torch.are_deterministic_algorithms_enabled()


def set_exact_backward(on) -> None:
    """True: every backward takes the deterministic kernel (gradients bit-identical to my_lib.cpp:91-130, ~5x the
    time of the atomic kernel, still ~0.1 % of a large forward).  False: always the atomic kernel (1e-5 contract).
    None (default): deterministic exactly when torch.use_deterministic_algorithms(True) is in force."""
    global _exact_backward
    _exact_backward = on


def exact_backward_enabled() -> bool:
    if _exact_backward is None:
        return bool(torch.are_deterministic_algorithms_enabled())
    return bool(_exact_backward)


def nnd_backward_raw(xyz1, xyz2, graddist1, graddist2, idx1, idx2, exact=None):
    """gradxyz1[B,N,3], gradxyz2[B,M,3] (my_lib.cpp:64-133 semantics, both always produced).
    exact=True: nnd_backward_exact_cuda_f32 (bit-identical to the reference CPU path); None: see set_exact_backward."""
    _check_clouds(xyz1, xyz2)
    lib = _cabi.load()
    xyz1 = xyz1.contiguous()
    xyz2 = xyz2.contiguous()
    b, n, _ = xyz1.shape
    m = xyz2.size(1)
    dev = xyz1.device
    if graddist1.dtype != torch.float32 or graddist1.device != dev:
        graddist1 = graddist1.to(device=dev, dtype=torch.float32)
    if graddist2.dtype != torch.float32 or graddist2.device != dev:
        graddist2 = graddist2.to(device=dev, dtype=torch.float32)
    graddist1 = graddist1.contiguous()  # __init__.py:40-41 (an expanded mean-gradient is materialised here)
    graddist2 = graddist2.contiguous()
    if graddist1.shape != (b, n) or graddist2.shape != (b, m):
        raise ValueError("graddist shapes must be [B,N] and [B,M]")
    if idx1.dtype != torch.int32 or idx2.dtype != torch.int32:
        raise TypeError("idx tensors must be int32 (__init__.py:19-20)")
    idx1 = idx1.contiguous()
    idx2 = idx2.contiguous()
    gradxyz1 = torch.empty_like(xyz1)
    gradxyz2 = torch.empty_like(xyz2)
    if exact is None:
        exact = exact_backward_enabled()
    with _on_device(dev):
        if exact:
            nbytes = lib.nnd_backward_exact_workspace_bytes(b, n, m)
            if nbytes == 0:
                raise ValueError("deterministic backward supports at most 2^31 - 1 points per cloud tensor")
            ws = _workspace(nbytes, dev)
            rc = lib.nnd_backward_exact_cuda_f32(xyz1.data_ptr(), xyz2.data_ptr(), graddist1.data_ptr(),
                                                 graddist2.data_ptr(), idx1.data_ptr(), idx2.data_ptr(), b, n, m,
                                                 gradxyz1.data_ptr(), gradxyz2.data_ptr(), ws.data_ptr(), ws.numel(),
                                                 _stream_ptr(dev))
            _cabi.check(rc, "nnd_backward_exact_cuda_f32")
            return gradxyz1, gradxyz2
        rc = lib.nnd_backward_cuda_f32(xyz1.data_ptr(), xyz2.data_ptr(), graddist1.data_ptr(), graddist2.data_ptr(),
                                       idx1.data_ptr(), idx2.data_ptr(), b, n, m, gradxyz1.data_ptr(),
                                       gradxyz2.data_ptr(), _stream_ptr(dev))
    if rc:
        _cabi.check(rc, "nnd_backward_cuda_f32")
    return gradxyz1, gradxyz2


def chamfer_reduce_workspace(device) -> torch.Tensor:
    """Scratch for chamfer_reduce_raw: uninitialised and stateless (the launcher prepares what it needs on the
    stream), so a fresh allocation per call is fine and nothing is shared between streams or graph captures."""
    lib = _cabi.load()
    return torch.empty(lib.nnd_chamfer_reduce_workspace_bytes(), dtype=torch.uint8, device=device)


def chamfer_reduce_raw(dist1, dist2, trunc: float, scale1: float, scale2: float, loss, graddist1, graddist2, ws):
    """loss[0] = scale1 * sum(dist1 where not >= trunc) + scale2 * (same for dist2); graddist = mask * scale.
    One launch (c2p-net/deformationpyramid/model/loss.py:150-215 with mean/mean reductions).  All outputs preallocated;
    graddist1/graddist2 may be None."""
    lib = _cabi.load()
    b, n = dist1.shape
    m = dist2.size(1)
    dev = dist1.device
    for t in (dist1, dist2, loss, graddist1, graddist2):
        if t is not None and (t.dtype != torch.float32 or not t.is_contiguous() or t.device != dev):
            raise ValueError("chamfer_reduce_raw wants contiguous float32 tensors on one device")
    with _on_device(dev):
        rc = lib.nnd_chamfer_reduce_cuda_f32(dist1.data_ptr(), dist2.data_ptr(), b, n, m, float(trunc), float(scale1),
                                             float(scale2), loss.data_ptr(),
                                             graddist1.data_ptr() if graddist1 is not None else None,
                                             graddist2.data_ptr() if graddist2 is not None else None, ws.data_ptr(),
                                             ws.numel(), _stream_ptr(dev))
    _cabi.check(rc, "nnd_chamfer_reduce_cuda_f32")


class NNDFunction(Function):
    """torch_nndistance.NNDFunction (reference __init__.py:10-57)."""

    @staticmethod
    def forward(ctx, xyz1, xyz2):
        dist1, dist2, idx1, idx2 = nnd_forward_raw(xyz1, xyz2)
        ctx.save_for_backward(xyz1, xyz2, dist1, dist2, idx1, idx2)  # __init__.py:34
        return dist1, dist2

    @staticmethod
    def backward(ctx, graddist1, graddist2):
        xyz1, xyz2, dist1, dist2, idx1, idx2 = ctx.saved_tensors
        return nnd_backward_raw(xyz1, xyz2, graddist1, graddist2, idx1, idx2)  # __init__.py:57


class _NNDWithIdxFunction(Function):
    @staticmethod
    def forward(ctx, xyz1, xyz2):
        dist1, dist2, idx1, idx2 = nnd_forward_raw(xyz1, xyz2)
        ctx.save_for_backward(xyz1, xyz2, idx1, idx2)
        ctx.mark_non_differentiable(idx1, idx2)
        return dist1, dist2, idx1, idx2

    @staticmethod
    def backward(ctx, graddist1, graddist2, _gi1, _gi2):
        xyz1, xyz2, idx1, idx2 = ctx.saved_tensors
        return nnd_backward_raw(xyz1, xyz2, graddist1, graddist2, idx1, idx2)


def nnd(xyz1, xyz2):
    """dist1, dist2 = nnd(xyz1[B,N,3], xyz2[B,M,3]) -- squared L2 to the nearest neighbour, both ways
    (reference __init__.py:60-61)."""
    return NNDFunction.apply(xyz1, xyz2)


def nnd_with_idx(xyz1, xyz2):
    """dist1, dist2, idx1, idx2 -- the same op with the int32 neighbour indices exposed (non-differentiable)."""
    return _NNDWithIdxFunction.apply(xyz1, xyz2)


class NNDModule(torch.nn.Module):
    """Legacy module name the reference smoke script still mentions (test.py:5-6)."""

    def forward(self, xyz1, xyz2):
        return nnd(xyz1, xyz2)
