"""Pins the caller-side rows to the REFERENCE'S OWN PYTHON, executed unmodified
(tests/ref_callers.py): c2p-net/deformationpyramid/model/loss.py:60-218, ROPNet/src/utils/process.py:14-27 +
ROPNet/src/loss/loss.py:52-59, c2p-net/deformationpyramid/model/geometry.py:412-452.

CPU tests (dev container, where /root/reference exists; skipped elsewhere):
  * the committed fixtures tests/golden/callers_*.npz are exactly what the reference code produces today;
  * the numpy restatement the GPU tests use (tests/test_compat.py::_cpu_chamfer) agrees with loss.py on the whole grid,
    NaN for an empty cloud included.
GPU tests (fixtures only, no /root/reference): pytorch3d_compat / callers on the B200 against those fixtures."""
import sys
from pathlib import Path

import numpy as np
import pytest
import torch

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
sys.path.insert(0, str(ROOT / "tests"))
sys.path.insert(0, str(ROOT / "tests" / "golden"))

import make_golden_callers as G  # noqa: E402
import ref_callers  # noqa: E402
from pointcloudregistration_b200 import synthetic  # noqa: E402

GOLDEN = ROOT / "tests" / "golden"
needs_ref = pytest.mark.skipif(not ref_callers.have_reference(), reason="/root/reference is not on this machine")


def _same(a, b):
    a, b = np.asarray(a), np.asarray(b)
    return a.shape == b.shape and np.array_equal(np.isnan(a), np.isnan(b)) and np.array_equal(np.nan_to_num(a), np.nan_to_num(b))


# ------------------------------------------------------------------------------------------ CPU, reference present
@needs_ref
def test_loss_fixture_is_what_the_reference_loss_py_produces():
    mod = ref_callers.load_loss_py()
    z = np.load(GOLDEN / "callers_loss.npz")
    for name, spec in G.LOSS_CASES.items():
        loss, gx, gy = G.run_loss_case(mod, spec)
        assert _same(loss, z[f"{name}.loss"]) and _same(gx, z[f"{name}.gx"]) and _same(gy, z[f"{name}.gy"]), name


@needs_ref
def test_numpy_restatement_of_the_chamfer_loss_agrees_with_the_reference_loss_py():
    from test_compat import _cpu_chamfer

    z = np.load(GOLDEN / "callers_loss.npz")
    x, y = G.loss_inputs()
    n, p1, p2 = x.shape[0], x.shape[1], y.shape[1]
    for name, (trunc, lx, ly, w, br, pr) in G.LOSS_CASES.items():
        lxa = np.full(n, p1) if lx is None else np.array(lx)
        lya = np.full(n, p2) if ly is None else np.array(ly)
        wa = None if w is None else np.array(w, np.float32)
        got = _cpu_chamfer(x.numpy(), y.numpy(), lxa, lya, None if np.isinf(trunc) else trunc, wa, br, pr)
        want = z[f"{name}.loss"]
        assert np.array_equal(np.isnan(got), np.isnan(want)), name
        np.testing.assert_allclose(np.nan_to_num(got), np.nan_to_num(want), rtol=2e-5, err_msg=name)


@needs_ref
def test_stub_knn_points_is_a_nearest_neighbour_search():
    x, y = synthetic.make_pair("crosscopy", 2, 90, 70, 83)
    nn = ref_callers.stub_knn_points(x, y, lengths1=torch.tensor([90, 11]), lengths2=torch.tensor([70, 3]))
    for b, (a, c) in enumerate(((90, 70), (11, 3))):
        D = ((x[b, :a, None, :] - y[b, None, :c, :]) ** 2).sum(-1)
        assert torch.equal(nn.idx[b, :a, 0], D.argmin(1)) and torch.allclose(nn.dists[b, :a, 0], D.min(1)[0], rtol=1e-6)
        assert (nn.dists[b, a:] == 0).all() and (nn.idx[b, a:] == 0).all()


@needs_ref
def test_ropnet_and_geometry_fixtures_are_what_the_reference_produces():
    square_dists, _ = ref_callers.load_ropnet()
    kind, b, n, m, seed = G.ROPNET_SHAPE
    src, tgt = synthetic.make_pair(kind, b, n, m, seed)
    dists = square_dists(src, tgt)
    z = np.load(GOLDEN / "callers_ropnet.npz")
    assert np.array_equal((torch.min(dists, dim=-1)[0] < 0.05 * 0.05).numpy().astype(np.int8), z["x_ol"])
    assert np.array_equal((torch.min(dists, dim=1)[0] < 0.05 * 0.05).numpy().astype(np.int8), z["y_ol"])
    geo = ref_callers.load_geometry()
    g = np.load(GOLDEN / "callers_geometry.npz")
    kind, b, n, m, seed = G.GEOM_CASES["uniform"]
    x, y = synthetic.make_pair(kind, b, n, m, seed)
    dist, idx = geo.knn_point_np(1, y[0].numpy(), x[0].numpy())
    assert np.array_equal(idx[:, 0], g["uniform.idx"]) and np.array_equal(dist[:, 0], g["uniform.dist"])


def test_oracle_agrees_with_the_reference_callers_fixtures():
    """CPU, fixtures only: the nnd oracle against the reference's ROPNet labels and geometry.py neighbours -- the
    measured disagreement DESIGN.md quotes (labels: 0 of 7600; KD-tree-free numpy knn: identical on tie-free input,
    ties resolved differently on lattice input because np.argpartition's order among equals is unspecified)."""
    import oracle

    z = np.load(GOLDEN / "callers_ropnet.npz")
    kind, b, n, m, seed = G.ROPNET_SHAPE
    src, tgt = synthetic.make_pair(kind, b, n, m, seed)
    d1, d2, _, _ = oracle.forward(src.numpy(), tgt.numpy(), threads=4)
    thr = 0.05 * 0.05
    differ = int(((d1 < thr) != z["x_ol"].astype(bool)).sum() + ((d2 < thr) != z["y_ol"].astype(bool)).sum())
    assert differ == 0
    assert max(np.abs(z["min_x"] - d1).max(), np.abs(z["min_y"] - d2).max()) < 2e-6  # the expansion form's rounding
    g = np.load(GOLDEN / "callers_geometry.npz")
    for name, (kind, b, n, m, seed) in G.GEOM_CASES.items():
        x, y = synthetic.make_pair(kind, b, n, m, seed)
        d, i = oracle.query(x.numpy(), y.numpy())
        assert np.array_equal(np.sqrt(d[0]), g[f"{name}.dist"]), name  # same distances bit for bit, ties or not
        if name == "uniform":
            assert np.array_equal(i[0], g[f"{name}.idx"])
        else:  # ties: the reference's pick is some index at the same distance; ours is the lowest such index
            assert (i[0] <= g[f"{name}.idx"]).all() and 0 < int((i[0] != g[f"{name}.idx"]).sum()) <= 5


# ------------------------------------------------------------------------------------------ GPU, fixtures only
@pytest.mark.gpu
@pytest.mark.parametrize("name", sorted(G.LOSS_CASES))
def test_compat_chamfer_matches_the_reference_loss_py(name, cuda_lib, device):
    """pytorch3d_compat.compute_truncated_chamfer_distance (and chamfer_distance where there is no truncation) on the
    B200 against what the reference's loss.py returned for the same inputs: loss to 2e-5 relative (the reference sums
    in fp32 torch order, the fused epilogue in float64), gradients to 1e-5 of the largest gradient, NaN where it has NaN."""
    from pointcloudregistration_b200 import pytorch3d_compat as P3D

    trunc, lx, ly, w, br, pr = G.LOSS_CASES[name]
    z = np.load(GOLDEN / "callers_loss.npz")
    x, y = G.loss_inputs()
    kw = dict(x_lengths=None if lx is None else torch.tensor(lx, device=device),
              y_lengths=None if ly is None else torch.tensor(ly, device=device),
              weights=None if w is None else torch.tensor(w, dtype=torch.float32, device=device),
              batch_reduction=br, point_reduction=pr)
    fns = [lambda a, c: P3D.compute_truncated_chamfer_distance(a, c, trunc=trunc, **kw)]
    if np.isinf(trunc):
        fns.append(lambda a, c: P3D.chamfer_distance(a, c, **kw)[0])
    for fn in fns:
        a = x.to(device).requires_grad_(True)
        c = y.to(device).requires_grad_(True)
        loss = fn(a, c)
        got, want = np.atleast_1d(loss.detach().cpu().numpy()), np.atleast_1d(z[f"{name}.loss"])
        assert np.array_equal(np.isnan(got), np.isnan(want)), f"{name}: NaN pattern {got} vs {want}"
        np.testing.assert_allclose(np.nan_to_num(got), np.nan_to_num(want), rtol=2e-5, atol=1e-12, err_msg=name)
        torch.where(torch.isfinite(loss), loss, torch.zeros_like(loss)).sum().backward()
        for g, wg, nm in ((a.grad, z[f"{name}.gx"], "gx"), (c.grad, z[f"{name}.gy"], "gy")):
            err = np.abs(g.cpu().numpy() - wg).max()
            assert err <= 1e-5 * max(np.abs(wg).max(), 1e-30), f"{name} {nm}: max abs err {err}"


@pytest.mark.gpu
def test_overlap_labels_match_the_reference_ropnet_labels(cuda_lib, device):
    from pointcloudregistration_b200 import callers

    z = np.load(GOLDEN / "callers_ropnet.npz")
    kind, b, n, m, seed = G.ROPNET_SHAPE
    src, tgt = synthetic.make_pair(kind, b, n, m, seed)
    xl, yl = callers.overlap_labels(src.to(device), tgt.to(device), dist_thresh=0.05)
    differ = int((xl.cpu().numpy() != z["x_ol"]).sum() + (yl.cpu().numpy() != z["y_ol"]).sum())
    assert differ == 0, f"{differ} of {z['x_ol'].size + z['y_ol'].size} labels differ from ROPNet's square_dists + min"


@pytest.mark.gpu
def test_numpy_helpers_match_the_reference_geometry_py(cuda_lib, device):
    from pointcloudregistration_b200 import callers

    g = np.load(GOLDEN / "callers_geometry.npz")
    for name, (kind, b, n, m, seed) in G.GEOM_CASES.items():
        x, y = synthetic.make_pair(kind, b, n, m, seed)
        src, tgt = x[0].numpy(), y[0].numpy()
        dist, idx = callers.nearest_neighbors_np(tgt, src)
        assert np.array_equal(dist, g[f"{name}.dist"]), name  # sqrt of bit-identical squared distances
        corr = callers.mutual_nn_correspondence(src, tgt, search_radius=0.05 if name == "uniform" else 0.3)
        if name == "uniform":
            assert np.array_equal(idx, g[f"{name}.idx"]) and np.array_equal(corr, g[f"{name}.corr"])
        else:  # exact ties: lowest index here, np.argpartition's unspecified pick there
            assert (idx <= g[f"{name}.idx"]).all()
            ours, ref = {tuple(c) for c in corr.T}, {tuple(c) for c in g[f"{name}.corr"].T}
            assert len(ours ^ ref) <= 6
