"""GPU tests of the thin caller helpers (pointcloudregistration_b200/callers.py) against numpy restatements of the
reference call sites they replace."""
import sys
from pathlib import Path

import numpy as np
import pytest
import torch

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))

from pointcloudregistration_b200 import callers, synthetic  # noqa: E402

pytestmark = pytest.mark.gpu


def _dense(a, b):
    d = b[None, :, :] - a[:, None, :]
    sq = d * d
    return (sq[..., 0] + sq[..., 1]) + sq[..., 2]  # [len(a), len(b)], the oracle's rounding order


def test_overlap_labels_match_dense_min(cuda_lib, device):
    x, y = synthetic.make_pair("clustered", 2, 600, 450, 90)
    xl, yl = callers.overlap_labels(x.to(device), y.to(device), dist_thresh=0.05)
    for b in range(2):
        D = _dense(x[b].numpy(), y[b].numpy())
        assert np.array_equal(xl[b].cpu().numpy(), (D.min(1) < 0.05 * 0.05).astype(np.int64))
        assert np.array_equal(yl[b].cpu().numpy(), (D.min(0) < 0.05 * 0.05).astype(np.int64))
    assert xl.dtype == torch.int64 and 0 < int(xl.sum()) < xl.numel()


def test_numpy_helpers(cuda_lib, device):
    x, y = synthetic.make_pair("uniform", 1, 500, 700, 91)
    a, b = x[0].numpy(), y[0].numpy()
    D = _dense(a, b)
    dist, idx = callers.nearest_neighbors_np(b, a)  # reference points first, like knn_point_np
    assert np.array_equal(idx, D.argmin(1).astype(np.int32))
    np.testing.assert_array_equal(dist, np.sqrt(D.min(1)))
    want = np.sqrt(D.min(1)).mean() + np.sqrt(D.min(0)).mean()
    assert abs(callers.chamfer_distance_np(a, b) - want) <= 1e-6 * want
    # mutual nearest neighbours (geometry.py:433-452)
    corr = callers.mutual_nn_correspondence(a, b, search_radius=0.05)
    s2t, t2s = D.argmin(1), D.argmin(0)
    keep = (t2s[s2t] == np.arange(len(a))) & (np.sqrt(D.min(1)) < 0.05)
    assert np.array_equal(corr, np.stack([np.arange(len(a))[keep], s2t[keep]]))
    with pytest.raises(ValueError):
        callers.nearest_neighbors_np(np.zeros((4, 2), np.float32), a)
