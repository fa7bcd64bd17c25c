"""Thin helpers for the reference's remaining 1-nearest-neighbour call sites,
all on top of the same B200 nnd kernels.  Each one names the reference site it can replace.

    overlap_labels            ROPNet/src/loss/loss.py:52-59, ROPNet/src/eval.py:59-77   (dense [B,N,M] square_dists + min)
    nearest_neighbors_np      dip/preprocess_lrf.py:101-110 (per-point open3d KD-tree search_knn_vector_xd(pt, 1)),
                              c2p-net/deformationpyramid/model/geometry.py:412-430 (knn_point_np with k == 1)
    chamfer_distance_np       DataPreparation/QualityCheck.py:25-31 (sklearn KD-tree, UNSQUARED distances, mean + mean)
    mutual_nn_correspondence  c2p-net/deformationpyramid/model/geometry.py:433-452

Numerics: distances are the exact separately-rounded fp32 squared differences of the nnd op.  ROPNet's
square_dists uses the |a|^2 + |b|^2 - 2ab expansion (ROPNet/src/utils/process.py:22-26), which rounds differently;
only a threshold is consumed there, so labels can differ for points within rounding of the threshold.  KD-tree sites
break exact ties by tree order; here ties go to the lowest index.  numpy in, numpy out; the arrays are moved to the
current CUDA device (there is no CPU path).
"""
from __future__ import annotations

import numpy as np
import torch

from .functional import nnd_forward_raw, nnd_query_raw


def overlap_labels(src: torch.Tensor, tgt: torch.Tensor, dist_thresh: float = 0.05):
    """x_ol_gt [B,N], y_ol_gt [B,M] int64: 1 where the nearest point of the other cloud is closer than dist_thresh
    (ROPNet/src/loss/loss.py:54-55).  O(N+M) memory instead of the dense [B,N,M] matrix."""
    with torch.no_grad():
        d1, d2, _, _ = nnd_forward_raw(src[..., :3].float().contiguous(), tgt[..., :3].float().contiguous())
    thr = dist_thresh * dist_thresh
    return (d1 < thr).long(), (d2 < thr).long()


def _to_device(a: np.ndarray, device) -> torch.Tensor:
    a = np.ascontiguousarray(a, dtype=np.float32)
    if a.ndim != 2 or a.shape[1] != 3:
        raise ValueError(f"expected an [N, 3] array, got {a.shape}")
    return torch.from_numpy(a).to(device)[None]


def _device(device):
    if device is not None:
        return torch.device(device)
    if not torch.cuda.is_available():
        raise RuntimeError("pointcloudregistration_b200 has no CPU path: a CUDA device is required")
    return torch.device("cuda", torch.cuda.current_device())


def nearest_neighbors_np(reference_pts: np.ndarray, query_pts: np.ndarray, device=None, squared: bool = False):
    """For every query point its nearest reference point: (dist [M] float32, idx [M] int32).
    geometry.py:412-430 returns sqrt distances (default here); pass squared=True for the raw op output."""
    dev = _device(device)
    d, i = nnd_query_raw(_to_device(query_pts, dev), _to_device(reference_pts, dev))
    d = d[0] if squared else torch.sqrt(d[0])
    return d.cpu().numpy(), i[0].cpu().numpy()


def chamfer_distance_np(p1: np.ndarray, p2: np.ndarray, device=None) -> float:
    """mean_x min_y |x - y| + mean_y min_x |x - y| with UNSQUARED distances (QualityCheck.py:25-31)."""
    dev = _device(device)
    d1, d2, _, _ = nnd_forward_raw(_to_device(p1, dev), _to_device(p2, dev))
    return float(torch.sqrt(d1).mean() + torch.sqrt(d2).mean())


def mutual_nn_correspondence(src_pcd_deformed: np.ndarray, tgt_pcd: np.ndarray, search_radius: float = 0.3,
                             knn: int = 1, device=None) -> np.ndarray:
    """[2, K] int array of (src index, tgt index) pairs that are each other's nearest neighbour and closer than
    search_radius (geometry.py:433-452).  One symmetric forward gives both directions."""
    if knn != 1:
        raise NotImplementedError("only knn == 1 (the value the reference uses)")
    dev = _device(device)
    d1, _, i1, i2 = nnd_forward_raw(_to_device(src_pcd_deformed, dev), _to_device(tgt_pcd, dev))
    ref_tgt_idx = i1[0].long()
    src_idx = torch.arange(ref_tgt_idx.numel(), device=dev)
    mutual = (i2[0].long()[ref_tgt_idx] == src_idx) & (torch.sqrt(d1[0]) < search_radius)
    return torch.stack([src_idx[mutual], ref_tgt_idx[mutual]], dim=0).cpu().numpy()
