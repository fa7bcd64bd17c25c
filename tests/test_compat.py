"""pytorch3d-shaped entry points (pointcloudregistration_b200/pytorch3d_compat.py): argument validation on
the CPU, results on the GPU against a numpy restatement built on the oracle's nnd.

Call sites being served: c2p-net/deformationpyramid/model/loss.py:60-218,143-144 and dip/train.py:84,113."""
import sys
from pathlib import Path

import numpy as np
import pytest
import torch

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
sys.path.insert(0, str(ROOT / "tests"))

from pointcloudregistration_b200 import pytorch3d_compat as P3D  # noqa: E402
from pointcloudregistration_b200 import synthetic  # noqa: E402


# ------------------------------------------------------------------------------------------ CPU
def test_unsupported_arguments_raise_before_any_cuda_work():
    a, b = torch.zeros(1, 4, 3), torch.zeros(1, 5, 3)
    with pytest.raises(NotImplementedError):
        P3D.knn_points(a, b, K=2)
    with pytest.raises(NotImplementedError):
        P3D.knn_points(torch.zeros(1, 4, 2), torch.zeros(1, 5, 2))
    with pytest.raises(ValueError):
        P3D.chamfer_distance(a, b, batch_reduction="max")
    with pytest.raises(ValueError):
        P3D.compute_truncated_chamfer_distance(a, b, point_reduction="median")
    with pytest.raises(ValueError):
        P3D.chamfer_distance(a, torch.zeros(2, 5, 3))
    with pytest.raises(ValueError):
        P3D.knn_points(a, b, lengths1=torch.tensor([1, 2]))
    with pytest.raises(RuntimeError):  # CPU tensors: there is no CPU path
        P3D.knn_points(a, b)


def test_knn_gather_matches_indexing():
    x = torch.arange(2 * 5 * 3, dtype=torch.float32).view(2, 5, 3)
    idx = torch.tensor([[[4], [0]], [[1], [1]]])
    out = P3D.knn_gather(x, idx)
    assert out.shape == (2, 2, 1, 3)
    assert torch.equal(out[0, 0, 0], x[0, 4]) and torch.equal(out[1, 1, 0], x[1, 1])


# ------------------------------------------------------------------------------------------ GPU
def _cpu_chamfer(x, y, lx, ly, trunc, weights, batch_reduction, point_reduction):
    """numpy restatement of loss.py:60-218 on top of the oracle's nnd (float64 reductions)."""
    import oracle

    n = x.shape[0]
    cx, cy = np.zeros((n, x.shape[1])), np.zeros((n, y.shape[1]))
    for i in range(n):
        a, b = int(lx[i]), int(ly[i])
        if a == 0 or b == 0:
            continue
        d1, d2, _, _ = oracle.forward(x[i:i + 1, :a], y[i:i + 1, :b])
        cx[i, :a], cy[i, :b] = d1[0], d2[0]
    if trunc is not None:
        cx[cx >= trunc] = 0.0
        cy[cy >= trunc] = 0.0
    if weights is not None:
        cx, cy = cx * weights[:, None], cy * weights[:, None]
    sx, sy = cx.sum(1), cy.sum(1)
    if point_reduction == "mean":
        with np.errstate(invalid="ignore", divide="ignore"):
            sx, sy = sx / lx, sy / ly  # loss.py:196-197 divides by the lengths as they are: empty cloud -> NaN
    if batch_reduction is not None:
        sx, sy = sx.sum(), sy.sum()
        if batch_reduction == "mean":
            div = weights.sum() if weights is not None else n
            sx, sy = sx / div, sy / div
    return sx + sy


@pytest.mark.gpu
def test_knn_points_matches_oracle_both_layouts(cuda_lib, device):
    import oracle

    x, y = synthetic.make_pair("crosscopy", 3, 700, 900, 80)
    nn = P3D.knn_points(x.to(device), y.to(device), K=1, return_nn=True)
    d, i = oracle.query(x.numpy(), y.numpy())
    assert nn.dists.shape == (3, 700, 1) and nn.idx.dtype == torch.int64 and nn.knn.shape == (3, 700, 1, 3)
    assert np.array_equal(nn.dists[..., 0].cpu().numpy(), d) and np.array_equal(nn.idx[..., 0].cpu().numpy(), i)
    assert torch.equal(nn.knn[:, :, 0].cpu(), torch.gather(y, 1, torch.from_numpy(i).long()[..., None].expand(-1, -1, 3)))
    # ragged: neighbours only among the first lengths2[i] points, padded query rows give 0
    l1, l2 = torch.tensor([700, 13, 0]), torch.tensor([900, 900, 5])
    nn = P3D.knn_points(x.to(device), y.to(device), lengths1=l1.to(device), lengths2=l2.to(device))
    for b in range(3):
        a, c = int(l1[b]), int(l2[b])
        got_d, got_i = nn.dists[b, :, 0].cpu().numpy(), nn.idx[b, :, 0].cpu().numpy()
        assert (got_d[a:] == 0).all() and (got_i[a:] == 0).all()
        if a:
            wd, wi = oracle.query(x[b:b + 1, :a].numpy(), y[b:b + 1, :c].numpy())
            assert np.array_equal(got_d[:a], wd[0]) and np.array_equal(got_i[:a], wi[0])


@pytest.mark.gpu
@pytest.mark.parametrize("trunc,ragged,weighted,br,pr", [
    (None, False, False, "mean", "mean"),      # dip/train.py:84
    (1e9, False, False, "mean", "mean"),       # registration.py:236 (trunc=1e9)
    (0.002, False, False, "mean", "mean"),     # real truncation
    (0.002, False, False, "sum", "mean"),      # fused epilogue, other scale combinations
    (None, False, False, "mean", "sum"),
    (0.002, True, True, "sum", "sum"),
    (None, True, False, None, "mean"),
])
def test_chamfer_matches_restatement_and_autograd(cuda_lib, device, trunc, ragged, weighted, br, pr):
    n, p1, p2 = 3, 400, 333
    x, y = synthetic.make_pair("clustered", n, p1, p2, 81)
    lx = np.array([400, 7, 250]) if ragged else np.full(n, p1)
    ly = np.array([333, 333, 1]) if ragged else np.full(n, p2)
    w = np.array([0.5, 2.0, 1.0], np.float32) if weighted else None
    a = x.to(device).requires_grad_(True)
    c = y.to(device).requires_grad_(True)
    kw = dict(x_lengths=torch.from_numpy(lx).to(device) if ragged else None,
              y_lengths=torch.from_numpy(ly).to(device) if ragged else None,
              weights=torch.from_numpy(w).to(device) if weighted else None, batch_reduction=br, point_reduction=pr)
    if trunc is None:
        got, normals = P3D.chamfer_distance(a, c, **kw)
        assert normals is None
    else:
        got = P3D.compute_truncated_chamfer_distance(a, c, trunc=trunc, **kw)
    want = _cpu_chamfer(x.numpy(), y.numpy(), lx, ly, trunc, w, br, pr)
    np.testing.assert_allclose(got.detach().cpu().numpy(), want, rtol=2e-5)
    got.sum().backward()
    assert a.grad.shape == x.shape and c.grad.shape == y.shape and torch.isfinite(a.grad).all()
    # padded rows never receive gradient
    for b in range(n):
        assert (a.grad[b, int(lx[b]):] == 0).all() and (c.grad[b, int(ly[b]):] == 0).all()
    # finite-difference check of one coordinate (the loss is piecewise quadratic: exact up to fp32 noise)
    b, j = 0, 5
    eps = 1e-3
    xp = x.clone()
    xp[b, j, 0] += eps
    xm = x.clone()
    xm[b, j, 0] -= eps
    f = lambda t: float(np.sum(_cpu_chamfer(t.numpy(), y.numpy(), lx, ly, trunc, w, br, pr)))  # noqa: E731
    fd = (f(xp) - f(xm)) / (2 * eps)
    assert abs(fd - float(a.grad[b, j, 0])) <= 5e-2 * max(abs(fd), 1e-6) + 1e-7


@pytest.mark.gpu
def test_chamfer_with_normals(cuda_lib, device):
    x, y = synthetic.make_pair("uniform", 2, 120, 90, 82)
    g = torch.Generator().manual_seed(5)
    xn = torch.nn.functional.normalize(torch.randn(2, 120, 3, generator=g), dim=2)
    yn = torch.nn.functional.normalize(torch.randn(2, 90, 3, generator=g), dim=2)
    loss, ln = P3D.chamfer_distance(x.to(device), y.to(device), x_normals=xn.to(device), y_normals=yn.to(device))
    import oracle

    _, _, i1, i2 = oracle.forward(x.numpy(), y.numpy())
    cx = 1 - np.abs((xn.numpy() * np.take_along_axis(yn.numpy(), i1[..., None].astype(np.int64), 1)).sum(-1))
    cy = 1 - np.abs((yn.numpy() * np.take_along_axis(xn.numpy(), i2[..., None].astype(np.int64), 1)).sum(-1))
    want = cx.mean(1).mean() + cy.mean(1).mean()
    np.testing.assert_allclose(float(ln), want, rtol=1e-4)
