"""Dev tool (GPU box): prepare / scan / finalize times (the library's own CUDA events, nnd_profile_collect) of one rank's
shard of the multi-GPU path.  usage: shard_phases.py M rows [rows ...]"""
import ctypes
import sys
from pathlib import Path

import torch

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from pointcloudregistration_b200 import _cabi  # noqa: E402
from pointcloudregistration_b200 import functional as F  # noqa: E402
from pointcloudregistration_b200 import synthetic  # noqa: E402

lib = _cabi.load()
m = int(sys.argv[1])
dev = torch.device("cuda:0")
x, y = synthetic.make_pair("uniform", 1, m, m, 4)
x, y = x.to(dev), y.to(dev)
for rows in [int(a) for a in sys.argv[2:]]:
    lo = (m // 3) // 8 * 8
    if lo + rows > m:
        lo = 0
    for _ in range(2):
        F.nnd_forward_rows_raw(x, y, lo, lo + rows)
    torch.cuda.synchronize()
    reps = 5
    lib.nnd_profile_enable(1)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        F.nnd_forward_rows_raw(x, y, lo, lo + rows)
    e1.record()
    torch.cuda.synchronize()
    ms = (ctypes.c_float * 3)()
    calls = ctypes.c_int(0)
    _cabi.check(lib.nnd_profile_collect(ms, ctypes.byref(calls)), "collect")
    lib.nnd_profile_enable(0)
    c = max(calls.value, 1)
    print(f"rows={rows:8d} x M={m}: total {e0.elapsed_time(e1) / reps:8.3f} ms | prepare {ms[0] / c:7.3f} | scan (+hints, seed) "
          f"{ms[1] / c:8.3f} | finalize {ms[2] / c:7.3f}  ({calls.value} profiled calls); ideal scan share "
          f"{rows / m:.4f} of the full problem", flush=True)
