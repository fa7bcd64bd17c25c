"""tests/golden/make_golden_callers.py -- fixtures for the caller-side rows, dev container only.

Executes the REFERENCE's own Python, unmodified (tests/ref_callers.py loads it from /root/reference):
  * c2p-net/deformationpyramid/model/loss.py:60-218  compute_truncated_chamfer_distance over a parameter grid
    (truncation, ragged lengths, weights, reductions, an empty cloud) -> loss values and gradients w.r.t. both clouds
  * ROPNet/src/utils/process.py:14-27 square_dists + ROPNet/src/loss/loss.py:54-55 overlap labels
  * c2p-net/deformationpyramid/model/geometry.py:412-452 knn_point_np / multual_nn_correspondence
and stores the outputs in tests/golden/callers_*.npz.  The GPU tests compare the B200 entry points
(pytorch3d_compat.py, callers.py) with these files on the GPU box, where /root/reference does not exist.

    python tests/golden/make_golden_callers.py
"""
from __future__ import annotations

import sys
from pathlib import Path

import numpy as np
import torch

ROOT = Path(__file__).resolve().parents[2]
sys.path.insert(0, str(ROOT))
sys.path.insert(0, str(ROOT / "tests"))

import ref_callers  # noqa: E402
from pointcloudregistration_b200 import synthetic  # noqa: E402

HERE = Path(__file__).resolve().parent

# name -> (trunc, x_lengths, y_lengths, weights, batch_reduction, point_reduction); clouds: clustered 3 x 400 / 333, seed 81
LOSS_CASES = {
    "dense_mean_mean_notrunc": (float("inf"), None, None, None, "mean", "mean"),   # dip/train.py:84 (chamfer_distance)
    "dense_mean_mean_trunc1e9": (1e9, None, None, None, "mean", "mean"),           # registration.py:236
    "dense_mean_mean_trunc": (0.002, None, None, None, "mean", "mean"),
    "dense_sum_mean_trunc": (0.002, None, None, None, "sum", "mean"),
    "dense_mean_sum_notrunc": (float("inf"), None, None, None, "mean", "sum"),
    "ragged_weighted_sum_sum_trunc": (0.002, [400, 7, 250], [333, 333, 1], [0.5, 2.0, 1.0], "sum", "sum"),
    "ragged_none_mean_notrunc": (float("inf"), [400, 7, 250], [333, 333, 1], None, None, "mean"),
    "ragged_mean_mean_trunc": (0.002, [400, 7, 250], [333, 333, 1], None, "mean", "mean"),
    "empty_cloud_sum_sum": (0.002, [400, 0, 250], [333, 333, 0], None, "sum", "sum"),
    "empty_cloud_none_mean": (0.002, [400, 0, 250], [333, 333, 5], None, None, "mean"),  # loss.py:196-197: 0/0 = NaN
}
LOSS_SHAPE = ("clustered", 3, 400, 333, 81)


def loss_inputs():
    kind, n, p1, p2, seed = LOSS_SHAPE
    return synthetic.make_pair(kind, n, p1, p2, seed)


def run_loss_case(mod, spec):
    trunc, lx, ly, w, br, pr = spec
    x, y = loss_inputs()
    x, y = x.clone().requires_grad_(True), y.clone().requires_grad_(True)
    kw = dict(x_lengths=None if lx is None else torch.tensor(lx), y_lengths=None if ly is None else torch.tensor(ly),
              weights=None if w is None else torch.tensor(w, dtype=torch.float32), trunc=trunc, batch_reduction=br,
              point_reduction=pr)
    loss = mod.compute_truncated_chamfer_distance(x, y, **kw)
    finite = torch.where(torch.isfinite(loss), loss, torch.zeros_like(loss))
    finite.sum().backward()  # gradients of the finite entries (a NaN entry has no usable gradient)
    return loss.detach().numpy(), x.grad.numpy(), y.grad.numpy()


def make_loss():
    mod = ref_callers.load_loss_py()
    out = {}
    for name, spec in LOSS_CASES.items():
        loss, gx, gy = run_loss_case(mod, spec)
        out[f"{name}.loss"], out[f"{name}.gx"], out[f"{name}.gy"] = loss, gx, gy
        print(f"loss.py {name}: loss = {np.array2string(np.atleast_1d(loss), precision=6)}")
    np.savez_compressed(HERE / "callers_loss.npz", **out)


ROPNET_SHAPE = ("uniform", 2, 2000, 1800, 90)


def make_ropnet():
    square_dists, Ol_loss = ref_callers.load_ropnet()
    kind, b, n, m, seed = ROPNET_SHAPE
    src, tgt = synthetic.make_pair(kind, b, n, m, seed)
    dists = square_dists(src, tgt)  # the |a|^2 + |b|^2 - 2ab expansion (process.py:22-26)
    x_ol = (torch.min(dists, dim=-1)[0] < 0.05 * 0.05).long()  # ROPNet/src/loss/loss.py:54
    y_ol = (torch.min(dists, dim=1)[0] < 0.05 * 0.05).long()   # ROPNet/src/loss/loss.py:55
    # the same two lines are what Ol_loss trains against: feed it logits that ARE the labels and check the loss is ~0
    big = 50.0
    lx = torch.stack([big * (1 - x_ol.float()), big * x_ol.float()], dim=1)  # [B, 2, N]
    ly = torch.stack([big * (1 - y_ol.float()), big * y_ol.float()], dim=1)
    assert float(Ol_loss(lx, ly, dists)) < 1e-6, "labels do not match the ones Ol_loss derives from dists"
    np.savez_compressed(HERE / "callers_ropnet.npz", x_ol=x_ol.numpy().astype(np.int8), y_ol=y_ol.numpy().astype(np.int8),
                        min_x=torch.min(dists, dim=-1)[0].numpy(), min_y=torch.min(dists, dim=1)[0].numpy())
    print(f"ROPNet labels: {int(x_ol.sum())} of {x_ol.numel()} src points, {int(y_ol.sum())} of {y_ol.numel()} tgt points overlap")


GEOM_CASES = {"uniform": ("uniform", 1, 500, 700, 91), "lattice": ("lattice", 1, 300, 300, 92)}


def make_geometry():
    geo = ref_callers.load_geometry()
    out = {}
    for name, (kind, b, n, m, seed) in GEOM_CASES.items():
        x, y = synthetic.make_pair(kind, b, n, m, seed)
        src, tgt = x[0].numpy(), y[0].numpy()
        dist, idx = geo.knn_point_np(1, tgt, src)  # reference points first (geometry.py:412)
        corr = geo.multual_nn_correspondence(src, tgt, search_radius=0.05 if kind == "uniform" else 0.3)
        out[f"{name}.dist"], out[f"{name}.idx"], out[f"{name}.corr"] = dist[:, 0], idx[:, 0].astype(np.int64), corr
        print(f"geometry.py {name}: {corr.shape[1]} mutual correspondences")
    np.savez_compressed(HERE / "callers_geometry.npz", **out)


if __name__ == "__main__":
    assert ref_callers.have_reference(), "/root/reference is needed (dev container)"
    make_loss()
    make_ropnet()
    make_geometry()
