"""Dev tool (GPU box): fraction of 4-target warp steps of nnd_scan_sym_kernel that enter the column slow path.
Needs a -DNND_SYM_COUNT build:  tools/variants.sh "count:-DNND_SYM_COUNT" ; NND_B200_DEBUG=1 NND_B200_LIB=tools/variants/libnnd_count.so"""
import ctypes
import math
import sys
from pathlib import Path

import torch

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from pointcloudregistration_b200 import _cabi, synthetic  # noqa: E402
from pointcloudregistration_b200 import functional as F  # noqa: E402

lib = _cabi.load()
raw = ctypes.CDLL(str(_cabi.LIB_PATH))
dev = torch.device("cuda:0")
out = (ctypes.c_ulonglong * 2)()
cases = [("C2", 32, 16384, 16384, None), ("C3", 1, 100_000, 100_000, None), ("C5", 8, 65536, 65536, None),
         ("C4/8 rows", 1, 1_000_000, 1_000_000, (0, 125_000)), ("C4", 1, 1_000_000, 1_000_000, None)]
for name, b, n, m, rows in cases:
    x, y = synthetic.make_pair("uniform", b, n, m, 5)
    x, y = x.to(dev), y.to(dev)
    raw.nnd_debug_sym_counters(out)
    if rows is None:
        F.nnd_forward_raw(x, y)
        nrows = n
    else:
        F.nnd_forward_rows_raw(x, y, *rows)
        nrows = rows[1] - rows[0]
    raw.nnd_debug_sym_counters(out)
    visits = nrows / 256  # warp visits per column (32 lanes x 8 rows)
    ideal = min(1.0, 4 * (math.log(visits) + 0.58) / visits)
    print(f"{name:10s} B={b} rows={nrows} M={m}: slow entries {out[0]:>12d} / steps {out[1]:>13d} = {out[0]/max(out[1],1):.4f}"
          f"   (sequential-filter estimate {ideal:.4f})", flush=True)
