"""pointcloudregistration_b200 -- B200-native (sm_100a) brute-force 3-D nearest
neighbour / chamfer op, a drop-in for the reference's dip/torch-nndistance.

Public surface (same names as the reference's torch_nndistance/__init__.py):
    nnd(xyz1, xyz2) -> (dist1, dist2)        NNDFunction        __version__
plus NNDModule, nnd_with_idx, set_exact_backward (bit-reproducible gradients; follows
torch.use_deterministic_algorithms by default), and the row-sharded multi-GPU entry points in
`pointcloudregistration_b200.sharded`.
"""
__version__ = "1.0.0"  # reference torch_nndistance/__init__.py:1

from .functional import (NNDFunction, NNDModule, nnd, nnd_backward_raw, nnd_forward_raw, nnd_query_raw,  # noqa: E402,F401
                         nnd_with_idx, set_exact_backward)
