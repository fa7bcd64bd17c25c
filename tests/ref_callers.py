"""TEST INFRASTRUCTURE: loads the reference's OWN Python for the caller-side rows so that it
can be executed, unmodified, as the checker -- in the dev container only (/root/reference does not exist on the GPU
box; the fixtures tests/golden/callers_*.npz made from it by tests/golden/make_golden_callers.py do travel).

    load_loss_py()      c2p-net/deformationpyramid/model/loss.py  (compute_truncated_chamfer_distance, :60-218)
                        -- it imports pytorch3d.ops.knn and pytorch3d.structures.pointclouds, which are neither vendored
                        nor installed (pin 276c9a8a..., c2p-net/requirements.txt:62); a stub of exactly those two modules
                        is injected: knn_points / knn_gather with pytorch3d's documented K = 1 semantics, the neighbour
                        search done by the nnd oracle (itself pinned bit-for-bit to the reference's my_lib.cpp)
    load_ropnet()       ROPNet/src/utils/process.py (square_dists, :14-27) and ROPNet/src/loss/loss.py (Ol_loss, :52-59)
    load_geometry()     c2p-net/deformationpyramid/model/geometry.py (knn_point_np :412-430, multual_nn_correspondence
                        :433-452); its `import open3d` is satisfied by an empty stub (those functions do not use it)
"""
from __future__ import annotations

import importlib.util
import sys
import types
from collections import namedtuple
from pathlib import Path

import numpy as np
import torch

REF = Path("/root/reference")
_KNN = namedtuple("KNN", "dists idx knn")


def have_reference() -> bool:
    return (REF / "c2p-net" / "deformationpyramid" / "model" / "loss.py").exists()


def _load(name: str, path: Path):
    spec = importlib.util.spec_from_file_location(name, str(path))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


# ------------------------------------------------------------------------------------------ pytorch3d stub (CPU)
def stub_knn_points(p1, p2, lengths1=None, lengths2=None, K=1, version=-1, return_nn=False, return_sorted=True):
    """pytorch3d.ops.knn_points, K = 1, on CPU tensors: dists [N,P1,1] squared L2 (0 for rows >= lengths1[i]),
    idx [N,P1,1] int64 (0 there).  The neighbour comes from the nnd oracle; the distance is differentiable w.r.t. both
    clouds (d = |p1 - p2[idx]|^2 through torch ops, value replaced by the oracle's bits)."""
    import oracle

    assert K == 1
    n, P1, _ = p1.shape
    P2 = p2.shape[1]
    l1 = [P1] * n if lengths1 is None else [int(v) for v in lengths1.tolist()]
    l2 = [P2] * n if lengths2 is None else [int(v) for v in lengths2.tolist()]
    idx = torch.zeros(n, P1, dtype=torch.int64)
    val = torch.zeros(n, P1, dtype=torch.float32)
    valid = torch.zeros(n, P1, dtype=torch.bool)
    a, c = p1.detach().numpy(), p2.detach().numpy()
    for i in range(n):
        if l1[i] == 0 or l2[i] == 0:
            continue
        d, j = oracle.query(a[i:i + 1, :l1[i]], c[i:i + 1, :l2[i]])
        idx[i, :l1[i]] = torch.from_numpy(j[0].astype(np.int64))
        val[i, :l1[i]] = torch.from_numpy(d[0])
        valid[i, :l1[i]] = True
    near = torch.gather(p2, 1, idx[..., None].expand(-1, -1, 3))
    formula = ((p1 - near) ** 2).sum(-1)
    dists = torch.where(valid, formula + (val - formula).detach(), torch.zeros_like(formula))
    nn = None
    if return_nn:
        nn = near[:, :, None, :]
    return _KNN(dists[..., None], idx[..., None], nn)


def stub_knn_gather(x, idx, lengths=None):
    """pytorch3d.ops.knn_gather: x [N,M,U], idx [N,L,K] -> [N,L,K,U]; entries with k >= lengths[i] are 0."""
    n, m, u = x.shape
    _, l, k = idx.shape
    out = x[:, :, None].expand(n, m, k, u).gather(1, idx[:, :, :, None].expand(n, l, k, u))
    if lengths is not None:
        bad = (torch.arange(k)[None] >= lengths[:, None]).view(n, 1, k, 1)
        out = out.masked_fill(bad, 0.0)
    return out


def _install_pytorch3d_stub():
    if "pytorch3d.ops.knn" in sys.modules and getattr(sys.modules["pytorch3d.ops.knn"], "_nnd_stub", False):
        return
    root, ops, knn = types.ModuleType("pytorch3d"), types.ModuleType("pytorch3d.ops"), types.ModuleType("pytorch3d.ops.knn")
    structures = types.ModuleType("pytorch3d.structures")
    pcl = types.ModuleType("pytorch3d.structures.pointclouds")

    class Pointclouds:  # loss.py only uses it in isinstance() checks
        pass

    knn.knn_points, knn.knn_gather, knn._nnd_stub = stub_knn_points, stub_knn_gather, True
    pcl.Pointclouds = Pointclouds
    root.ops, ops.knn, root.structures, structures.pointclouds = ops, knn, structures, pcl
    sys.modules.update({"pytorch3d": root, "pytorch3d.ops": ops, "pytorch3d.ops.knn": knn,
                        "pytorch3d.structures": structures, "pytorch3d.structures.pointclouds": pcl})


def load_loss_py():
    """The reference's loss.py, unmodified, on top of the stub above."""
    _install_pytorch3d_stub()
    return _load("_ref_c2p_loss", REF / "c2p-net" / "deformationpyramid" / "model" / "loss.py")


def load_ropnet():
    """(square_dists, Ol_loss) of the reference's ROPNet."""
    process = _load("_ref_ropnet_process", REF / "ROPNet" / "src" / "utils" / "process.py")
    shim = types.ModuleType("utils")  # ROPNet/src/loss/loss.py:4 does `from utils import square_dists`
    shim.square_dists = process.square_dists
    saved = sys.modules.get("utils")
    sys.modules["utils"] = shim
    try:
        loss = _load("_ref_ropnet_loss", REF / "ROPNet" / "src" / "loss" / "loss.py")
    finally:
        if saved is None:
            sys.modules.pop("utils", None)
        else:
            sys.modules["utils"] = saved
    return process.square_dists, loss.Ol_loss


def load_geometry():
    """The reference's geometry.py (knn_point_np, multual_nn_correspondence)."""
    if "open3d" not in sys.modules:
        sys.modules["open3d"] = types.ModuleType("open3d")  # imported at geometry.py:4, unused by these functions
    return _load("_ref_c2p_geometry", REF / "c2p-net" / "deformationpyramid" / "model" / "geometry.py")
