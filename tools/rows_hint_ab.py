"""Dev tool (GPU box): one rank's shard of the multi-GPU path (nnd_forward_rows on a row range) with the automatic
hints off / on, uniform and clustered clouds.  usage: rows_hint_ab.py M rows [rows ...]"""
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
keep = lib.nnd_set_auto_hint(-1)
for kind in ("uniform", "clustered"):
    x, y = synthetic.make_pair(kind, 1, m, m, 4)
    x, y = x.to(dev), y.to(dev)
    for rows in [int(a) for a in sys.argv[2:]]:
        lo = (m // 3) // 8 * 8
        if lo + rows > m:
            lo = 0
        line = f"{kind:9s} rows={rows:8d} x M={m} (scan kind {lib.nnd_rows_scan_kind(1, m, m, lo, lo + rows)}):"
        for mode in (0, 1):
            lib.nnd_set_auto_hint(mode)
            for _ in range(2):
                F.nnd_forward_rows_raw(x, y, lo, lo + rows)
            torch.cuda.synchronize()
            ts = []
            for _ in range(5):
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record()
                F.nnd_forward_rows_raw(x, y, lo, lo + rows)
                e1.record()
                torch.cuda.synchronize()
                ts.append(e0.elapsed_time(e1))
            t = min(ts)
            line += f" | hints {mode}: {t:9.3f} ms {rows * m / t / 1e9:6.3f} Tevals/s"
        print(line, flush=True)
lib.nnd_set_auto_hint(keep)
