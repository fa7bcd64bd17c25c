"""Drop-in `torch_nndistance_aten` module name (the pybind module the reference builds,
dip/torch-nndistance/build.py:52; exports my_lib_cuda.cpp:74-77 / my_lib.cpp:135-138)."""
from pointcloudregistration_b200.aten import (nnd_backward, nnd_backward_cuda, nnd_forward, nnd_forward_cuda)  # noqa: F401
