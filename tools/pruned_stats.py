"""Dev tool (GPU box): work counters of the pruned search from an instrumented build (-DNND_PR_STATS):
NND_B200_DEBUG=1 NND_B200_LIB=tools/variants/libnnd_prstats.so python tools/pruned_stats.py"""
import ctypes
import sys
from pathlib import Path

import torch

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from pointcloudregistration_b200 import _cabi  # noqa: E402
from pointcloudregistration_b200 import functional as F  # noqa: E402
from pointcloudregistration_b200 import synthetic  # noqa: E402

lib = _cabi.load()
dev = torch.device("cuda:0")
out = (ctypes.c_ulonglong * 4)()
for name, kind, b, n, cfg in (("C4 uniform", "uniform", 1, 1_000_000, 4), ("C4 clustered", "clustered", 1, 1_000_000, 4),
                              ("C5 lattice", "lattice", 1, 262_144, 5), ("C3 uniform", "uniform", 1, 100_000, 3),
                              ("C2 uniform", "uniform", 32, 16384, 2)):
    x1, x2 = synthetic.make_pair(kind, b, n, n, cfg)
    a, c = x1.to(dev), x2.to(dev)
    lib.nnd_debug_pruned_stats(out)
    F.nnd_forward_pruned_raw(a, c)
    torch.cuda.synchronize()
    lib.nnd_debug_pruned_stats(out)
    scan, sup, til, warps = [int(v) for v in out]
    print(f"{name:14s}: warps {warps}, per warp: tiles scanned {scan / warps:7.1f}, super-tiles passed {sup / warps:6.1f}, "
          f"tiles passed the box test {til / warps:7.1f}", flush=True)
