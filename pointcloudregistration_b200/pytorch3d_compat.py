"""pytorch3d-shaped entry points backed by the B200 nnd kernels.

This fork of the reference rewired its callers from torch_nndistance to pytorch3d:
    c2p-net/deformationpyramid/model/loss.py:4,143-144   knn_points(x, y, lengths1, lengths2, K=1) both ways
    c2p-net/deformationpyramid/model/loss.py:60-218      compute_truncated_chamfer_distance
    dip/train.py:13,84,113                               chamfer_distance(x, y) -> (loss, None)
These functions keep those call sites working with a one-line import change.  Same names, argument
meaning and return shapes; only what the call sites use is implemented (K == 1, D == 3, squared L2):
anything else raises NotImplementedError instead of silently computing something different.

pytorch3d (pin 276c9a8a..., c2p-net/requirements.txt:62) is not vendored in the reference and not
installed here, so parity at THIS boundary is unpinned; the arithmetic underneath is the nnd op, which is
pinned bit-exactly against the reference CPU path.  Documented behaviour followed:
dists are squared distances, idx is int64, outputs of padded query rows (row >= lengths1[i]) are 0,
neighbours are searched among the first lengths2[i] points only.

CUDA tensors only: like the rest of the package there is no CPU path.
"""
from __future__ import annotations

from collections import namedtuple
from typing import Optional, Union

import torch
from torch.autograd import Function

from . import _cabi
from . import functional as F_
from .functional import _check_clouds, _stream_ptr, nnd_backward_raw, nnd_forward_raw, nnd_query_raw

_KNN = namedtuple("KNN", "dists idx knn")


def _lengths(points: torch.Tensor, lengths: Optional[torch.Tensor], name: str) -> Optional[torch.Tensor]:
    """The per-cloud lengths as given (a tensor [N]) or None.  No host read here: shapes that take the single-launch
    kernels hand the tensor to the kernel as it is (the kernel clamps every length to [0, P])."""
    if lengths is None:
        return None
    if lengths.ndim != 1 or lengths.shape[0] != points.shape[0]:
        raise ValueError(f"Expected {name} to be of shape (N,)")  # loss.py:44-47
    return lengths


def _length_list(points: torch.Tensor, lengths: Optional[torch.Tensor], name: str) -> Optional[list]:
    """Host copy of the lengths for the per-cloud loop (large ragged clouds only): None when every cloud is full."""
    if lengths is None:
        return None
    vals = [int(v) for v in lengths.tolist()]
    if any(v < 0 or v > points.shape[1] for v in vals):
        raise ValueError(f"{name} must lie in [0, P]")
    return None if all(v == points.shape[1] for v in vals) else vals


def _backward_term(q, t, gd, idx, grad_q, grad_t) -> None:
    """grad_q[row] += 2*gd*(q - t[idx]); grad_t[idx] -= the same (include/nnd_b200.h, accumulating)."""
    lib = _cabi.load()
    b, nq, _ = q.shape
    if b == 0 or nq == 0 or t.size(1) == 0:
        return
    with torch.cuda.device(q.device):
        rc = lib.nnd_backward_term_cuda_f32(q.data_ptr(), t.data_ptr(), gd.data_ptr(), idx.data_ptr(), b, nq,
                                            t.size(1), grad_q.data_ptr(), grad_t.data_ptr(), _stream_ptr(q.device))
    _cabi.check(rc, "nnd_backward_term_cuda_f32")


def _one_launch(p1, p2) -> bool:
    """Ragged batches of this shape are handled by ONE launch for the whole batch (nnd_*_ragged_cuda_f32)."""
    return p1.size(0) > 0 and p1.size(1) > 0 and p2.size(1) > 0 and F_.ragged_supported(p1.size(1), p2.size(1))


class _KNN1Function(Function):
    """One direction, K = 1: dists [N,P1] f32 (squared), idx [N,P1] int32; rows >= lengths1[i] are 0."""

    @staticmethod
    def forward(ctx, p1, p2, len1, len2):
        _check_clouds(p1, p2)
        p1c, p2c = p1.contiguous(), p2.contiguous()
        ctx.one_launch = False
        if len1 is None and len2 is None:
            dists, idx = nnd_query_raw(p1c, p2c)
        elif _one_launch(p1c, p2c):
            ctx.one_launch = True
            dists, idx = F_.nnd_query_ragged_raw(p1c, p2c, len1, len2)
        else:
            len1, len2 = _length_list(p1c, len1, "lengths1"), _length_list(p2c, len2, "lengths2")
            n = p1c.size(0)
            dists = p1c.new_zeros(n, p1c.size(1))
            idx = torch.zeros(n, p1c.size(1), dtype=torch.int32, device=p1c.device)
            for i in range(n):  # large ragged clouds: one forward per cloud on its valid prefix
                l1 = p1c.size(1) if len1 is None else len1[i]
                l2 = p2c.size(1) if len2 is None else len2[i]
                if l1 == 0 or l2 == 0:
                    continue
                d, j = nnd_query_raw(p1c[i:i + 1, :l1], p2c[i:i + 1, :l2])
                dists[i, :l1], idx[i, :l1] = d[0], j[0]
        ctx.save_for_backward(p1c, p2c, idx)
        ctx.len1, ctx.len2 = len1, len2
        ctx.mark_non_differentiable(idx)
        return dists, idx

    @staticmethod
    def backward(ctx, gdists, _gidx):
        p1, p2, idx = ctx.saved_tensors
        len1, len2 = ctx.len1, ctx.len2
        gdists = gdists.to(torch.float32).contiguous()
        if ctx.one_launch:
            g1, g2 = F_.nnd_knn_backward_ragged_raw(p1, p2, gdists, idx, len1, len2)
            return g1, g2, None, None
        g1, g2 = torch.zeros_like(p1), torch.zeros_like(p2)
        if len1 is None and len2 is None:
            _backward_term(p1, p2, gdists, idx, g1, g2)
        else:
            for i in range(p1.size(0)):
                l1 = p1.size(1) if len1 is None else len1[i]
                l2 = p2.size(1) if len2 is None else len2[i]
                if l1 == 0 or l2 == 0:
                    continue
                q, t = p1[i:i + 1, :l1].contiguous(), p2[i:i + 1, :l2].contiguous()
                gq, gt = torch.zeros_like(q), torch.zeros_like(t)
                _backward_term(q, t, gdists[i:i + 1, :l1].contiguous(), idx[i:i + 1, :l1].contiguous(), gq, gt)
                g1[i, :l1], g2[i, :l2] = gq[0], gt[0]
        return g1, g2, None, None


class _NND2Function(Function):
    """Both directions, K = 1, with optional per-cloud lengths.  The homogeneous case is ONE symmetric
    forward (every distance evaluated once, csrc/nnd_sym.cuh) instead of pytorch3d's two knn passes; ragged batches
    of small clouds are one launch too (per-cloud lengths read by the kernel, no host synchronisation)."""

    @staticmethod
    def forward(ctx, x, y, lx, ly):
        _check_clouds(x, y)
        xc, yc = x.contiguous(), y.contiguous()
        ctx.one_launch = False
        if lx is None and ly is None:
            d1, d2, i1, i2 = nnd_forward_raw(xc, yc)
        elif _one_launch(xc, yc):
            ctx.one_launch = True
            d1, d2, i1, i2 = F_.nnd_forward_ragged_raw(xc, yc, lx, ly)
        else:
            lx, ly = _length_list(xc, lx, "x_lengths"), _length_list(yc, ly, "y_lengths")
            n = xc.size(0)
            d1, d2 = xc.new_zeros(n, xc.size(1)), yc.new_zeros(n, yc.size(1))
            i1 = torch.zeros(n, xc.size(1), dtype=torch.int32, device=xc.device)
            i2 = torch.zeros(n, yc.size(1), dtype=torch.int32, device=xc.device)
            for i in range(n):
                a = xc.size(1) if lx is None else lx[i]
                b = yc.size(1) if ly is None else ly[i]
                if a == 0 or b == 0:
                    continue
                e1, e2, j1, j2 = nnd_forward_raw(xc[i:i + 1, :a], yc[i:i + 1, :b])
                d1[i, :a], d2[i, :b], i1[i, :a], i2[i, :b] = e1[0], e2[0], j1[0], j2[0]
        ctx.save_for_backward(xc, yc, i1, i2)
        ctx.lx, ctx.ly = lx, ly
        ctx.mark_non_differentiable(i1, i2)
        return d1, d2, i1, i2

    @staticmethod
    def backward(ctx, g1, g2, _a, _b):
        x, y, i1, i2 = ctx.saved_tensors
        lx, ly = ctx.lx, ctx.ly
        g1, g2 = g1.to(torch.float32).contiguous(), g2.to(torch.float32).contiguous()
        if ctx.one_launch:
            gx, gy = F_.nnd_backward_ragged_raw(x, y, g1, g2, i1, i2, lx, ly)
            return gx, gy, None, None
        if lx is None and ly is None:
            gx, gy = nnd_backward_raw(x, y, g1, g2, i1, i2)
            return gx, gy, None, None
        gx, gy = torch.zeros_like(x), torch.zeros_like(y)
        for i in range(x.size(0)):
            a = x.size(1) if lx is None else lx[i]
            b = y.size(1) if ly is None else ly[i]
            if a == 0 or b == 0:
                continue
            ex, ey = nnd_backward_raw(x[i:i + 1, :a], y[i:i + 1, :b], g1[i:i + 1, :a], g2[i:i + 1, :b],
                                      i1[i:i + 1, :a].contiguous(), i2[i:i + 1, :b].contiguous())
            gx[i, :a], gy[i, :b] = ex[0], ey[0]
        return gx, gy, None, None


def _check_k_d(p1, p2, K):
    if K != 1:
        raise NotImplementedError("only K == 1 is implemented (the only value the reference's call sites use)")
    if p1.shape[-1] != 3 or p2.shape[-1] != 3:
        raise NotImplementedError("only D == 3 is implemented")


def knn_points(p1, p2, lengths1=None, lengths2=None, K: int = 1, version: int = -1, return_nn: bool = False,
               return_sorted: bool = True):
    """pytorch3d.ops.knn_points for K = 1, D = 3 (call sites: loss.py:143-144).

    Returns KNN(dists [N,P1,1] f32 squared L2, idx [N,P1,1] int64, knn [N,P1,1,3] or None).
    Differentiable w.r.t. p1 and p2 through dists (and through knn when return_nn)."""
    _check_k_d(p1, p2, K)
    l1, l2 = _lengths(p1, lengths1, "lengths1"), _lengths(p2, lengths2, "lengths2")
    dists, idx = _KNN1Function.apply(p1, p2, l1, l2)
    idx64 = idx.to(torch.int64).unsqueeze(-1)
    nn = knn_gather(p2, idx64, lengths2) if return_nn else None
    return _KNN(dists.unsqueeze(-1), idx64, nn)


def knn_gather(x, idx, lengths=None):
    """pytorch3d.ops.knn_gather: x [N,M,U], idx [N,L,K] -> [N,L,K,U]; neighbours beyond lengths[i] give 0
    (used by loss.py:168-169 for the normals)."""
    n, m, u = x.shape
    _, l, k = idx.shape
    out = x[:, :, None].expand(n, m, k, u).gather(1, idx[:, :, :, None].expand(n, l, k, u))
    if lengths is not None:
        bad = (torch.arange(k, device=x.device)[None] >= lengths[:, None]).view(n, 1, k, 1)
        out = out.masked_fill(bad, 0.0)  # fewer than K valid neighbours
    return out


def _nnd2(x, y, x_lengths, y_lengths):
    lx, ly = _lengths(x, x_lengths, "x_lengths"), _lengths(y, y_lengths, "y_lengths")
    return _NND2Function.apply(x, y, lx, ly)


def _reduce(cham_x, cham_y, x_lengths, y_lengths, weights, batch_reduction, point_reduction, n):
    """Point and batch reductions of loss.py:185-215 (identical in pytorch3d's chamfer_distance)."""
    cham_x, cham_y = cham_x.sum(1), cham_y.sum(1)
    if point_reduction == "mean":
        cham_x = cham_x / x_lengths  # loss.py:196-197: no clamp -- an empty cloud gives 0/0 = NaN there and here
        cham_y = cham_y / y_lengths
    if batch_reduction is not None:
        cham_x, cham_y = cham_x.sum(), cham_y.sum()
        if batch_reduction == "mean":
            div = weights.sum() if weights is not None else n  # loss.py:210
            cham_x, cham_y = cham_x / div, cham_y / div
    return cham_x + cham_y


def _validate(batch_reduction, point_reduction):
    if batch_reduction is not None and batch_reduction not in ["mean", "sum"]:
        raise ValueError('batch_reduction must be one of ["mean", "sum"] or None')  # loss.py:18-19
    if point_reduction not in ["mean", "sum"]:
        raise ValueError('point_reduction must be one of ["mean", "sum"]')  # loss.py:20-21


def _full_lengths(points, lengths):
    if lengths is None:
        return torch.full((points.shape[0],), points.shape[1], dtype=torch.int64, device=points.device)
    return lengths.to(device=points.device, dtype=torch.int64)


class _FusedChamferFunction(Function):
    """Dense, unweighted chamfer with scalar output: nnd forward + ONE fused launch for mask, both reductions and
    d loss/d dist (nnd_chamfer_reduce_cuda_f32); the backward scales the saved d loss/d dist and runs nnd's backward.
    Same arithmetic as the general path below up to the summation order (float64 partial sums here)."""

    @staticmethod
    def forward(ctx, x, y, trunc, s1, s2):
        xc, yc = x.detach().contiguous(), y.detach().contiguous()  # dtype/device checked by nnd_forward_raw
        d1, d2, i1, i2 = F_.nnd_forward_raw(xc, yc)
        ws = F_.chamfer_reduce_workspace(xc.device)  # per call: stateless scratch, nothing cached across streams
        loss = torch.empty((), dtype=torch.float32, device=xc.device)
        gd1, gd2 = torch.empty_like(d1), torch.empty_like(d2)
        F_.chamfer_reduce_raw(d1, d2, trunc, s1, s2, loss, gd1, gd2, ws)
        ctx.save_for_backward(xc, yc, gd1, gd2, i1, i2)
        return loss

    @staticmethod
    def backward(ctx, g):
        xc, yc, gd1, gd2, i1, i2 = ctx.saved_tensors
        gx, gy = F_.nnd_backward_raw(xc, yc, gd1 * g, gd2 * g, i1, i2)
        return gx, gy, None, None, None


def _chamfer(x, y, x_lengths, y_lengths, x_normals, y_normals, weights, trunc, batch_reduction, point_reduction):
    _validate(batch_reduction, point_reduction)
    if x.ndim != 3 or y.ndim != 3:
        raise ValueError("Expected points to be of shape (N, P, D)")  # loss.py:40-41
    n, p1, d = x.shape
    p2 = y.shape[1]
    if y.shape[0] != n or y.shape[2] != d:
        raise ValueError("y does not have the correct shape.")  # loss.py:121-122
    if weights is not None:
        if weights.size(0) != n:
            raise ValueError("weights must be of shape (N,).")
        if not (weights >= 0).all():
            raise ValueError("weights cannot be negative.")
        if weights.sum() == 0.0:  # loss.py:128-135
            z = (x.sum((1, 2)) * weights.view(n)) * 0.0
            return (z.sum(), z.sum()) if batch_reduction in ["mean", "sum"] else (z, z)
    if (weights is None and x_lengths is None and y_lengths is None and batch_reduction is not None
            and (x_normals is None or y_normals is None) and n > 0 and p1 > 0 and p2 > 0 and d == 3):
        _check_k_d(x, y, 1)
        bdiv = n if batch_reduction == "mean" else 1
        s1 = 1.0 / ((p1 if point_reduction == "mean" else 1) * bdiv)
        s2 = 1.0 / ((p2 if point_reduction == "mean" else 1) * bdiv)
        # no truncation = a NaN threshold: `dist >= NaN` is never true, so nothing is dropped (not even +inf)
        return _FusedChamferFunction.apply(x, y, float("nan") if trunc is None else float(trunc), s1, s2), None
    xl, yl = _full_lengths(x, x_lengths), _full_lengths(y, y_lengths)
    cham_x, cham_y, idx_x, idx_y = _nnd2(x, y, x_lengths, y_lengths)
    x_mask = torch.arange(p1, device=x.device)[None] >= xl[:, None]  # loss.py:113-118
    y_mask = torch.arange(p2, device=y.device)[None] >= yl[:, None]
    if trunc is not None:  # loss.py:151-154: distances >= trunc do not count
        x_mask = x_mask | (cham_x.detach() >= trunc)
        y_mask = y_mask | (cham_y.detach() >= trunc)
    cham_x = cham_x.masked_fill(x_mask, 0.0)
    cham_y = cham_y.masked_fill(y_mask, 0.0)
    if weights is not None:
        cham_x = cham_x * weights.view(n, 1)
        cham_y = cham_y * weights.view(n, 1)
    dist = _reduce(cham_x, cham_y, xl, yl, weights, batch_reduction, point_reduction, n)
    normals = None
    if x_normals is not None and y_normals is not None:  # loss.py:166-183
        import torch.nn.functional as F

        xn_near = knn_gather(y_normals, idx_x.to(torch.int64).unsqueeze(-1), y_lengths)[..., 0, :]
        yn_near = knn_gather(x_normals, idx_y.to(torch.int64).unsqueeze(-1), x_lengths)[..., 0, :]
        cnx = 1 - torch.abs(F.cosine_similarity(x_normals, xn_near, dim=2, eps=1e-6))
        cny = 1 - torch.abs(F.cosine_similarity(y_normals, yn_near, dim=2, eps=1e-6))
        pad_x = torch.arange(p1, device=x.device)[None] >= xl[:, None]
        pad_y = torch.arange(p2, device=y.device)[None] >= yl[:, None]
        cnx, cny = cnx.masked_fill(pad_x, 0.0), cny.masked_fill(pad_y, 0.0)
        if weights is not None:
            cnx, cny = cnx * weights.view(n, 1), cny * weights.view(n, 1)
        normals = _reduce(cnx, cny, xl, yl, weights, batch_reduction, point_reduction, n)
    return dist, normals


def chamfer_distance(x, y, x_lengths=None, y_lengths=None, x_normals=None, y_normals=None, weights=None,
                     batch_reduction: Union[str, None] = "mean", point_reduction: str = "mean"):
    """pytorch3d.loss.chamfer_distance as dip/train.py:84,113 calls it: (loss, loss_normals or None),
    squared L2, sum of both directions, mean over points then over the batch by default."""
    return _chamfer(x, y, x_lengths, y_lengths, x_normals, y_normals, weights, None, batch_reduction,
                    point_reduction)


def compute_truncated_chamfer_distance(x, y, x_lengths=None, y_lengths=None, x_normals=None, y_normals=None,
                                       weights=None, trunc=0.2, batch_reduction: Union[str, None] = "mean",
                                       point_reduction: str = "mean"):
    """c2p-net/deformationpyramid/model/loss.py:60-218: chamfer where nearest-neighbour distances >= trunc are
    dropped; returns the distance term only, like the reference (loss.py:218)."""
    return _chamfer(x, y, x_lengths, y_lengths, x_normals, y_normals, weights, trunc, batch_reduction,
                    point_reduction)[0]
