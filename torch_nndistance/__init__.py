"""Drop-in `torch_nndistance` package name (reference dip/torch-nndistance/torch_nndistance/__init__.py).

`import torch_nndistance as NND; dist1, dist2 = NND.nnd(xyz1, xyz2)` keeps working;
the implementation is pointcloudregistration_b200 (hand-written sm_100a kernels).
"""
from pointcloudregistration_b200 import (NNDFunction, NNDModule, __version__, nnd, nnd_with_idx)  # noqa: F401
