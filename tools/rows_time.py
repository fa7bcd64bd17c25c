"""Dev tool (GPU box): time nnd_forward_rows (one rank's shard of the multi-GPU path) for a few row counts.
usage: rows_time.py M rows [rows ...]"""
import sys
from pathlib import Path

import torch

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from pointcloudregistration_b200 import functional as F  # noqa: E402
from pointcloudregistration_b200 import synthetic  # noqa: E402

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
    ts = []
    for _ in range(4):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        F.nnd_forward_rows_raw(x, y, lo, lo + rows)
        e1.record()
        torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    t = sorted(ts)[len(ts) // 2]
    print(f"rows={rows:8d} x M={m}: {t:9.3f} ms   {rows * m / t / 1e9:7.2f} Tevals/s", flush=True)
