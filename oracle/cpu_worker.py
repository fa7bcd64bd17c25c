"""oracle/cpu_worker.py -- TEST / BENCH INFRASTRUCTURE ONLY (one CPU worker of bench.py's CPU arm).

Runs the reference's own nnd_forward + nnd_backward (oracle/_ref, my_lib.cpp compiled unmodified;
`--kind port` uses the C restatement instead) on one slice [lo, hi) of cloud-1 query rows against
the full cloud 2, `warmup + steps` times, starting at wall-clock time --start-at so that all workers
load the host cores together.  Prints one JSON line: {"kind": ..., "step_seconds": [...]}.
"""
from __future__ import annotations

import argparse
import json
import sys
import time
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--kind", choices=["reference", "port"], default="reference")
    for name in ("b", "n", "m", "config", "lo", "hi", "steps", "warmup"):
        ap.add_argument(f"--{name}", type=int, required=True)
    ap.add_argument("--cloud", default="uniform")
    ap.add_argument("--start-at", type=float, default=0.0)
    a = ap.parse_args()

    import numpy as np
    import torch

    torch.set_num_threads(1)
    import oracle
    from pointcloudregistration_b200 import synthetic

    x1, x2 = synthetic.make_pair(a.cloud, a.b, a.n, a.m, a.config)
    lo, hi = max(0, a.lo), min(a.n, a.hi)
    q = x1[:, lo:hi].contiguous()
    gd1 = torch.full((a.b, hi - lo), 1.0 / (a.b * a.n))
    gd2 = torch.full((a.b, a.m), 1.0 / (a.b * a.m))
    kind = a.kind
    if kind == "reference" and not oracle.have_reference():
        kind = "port"
    if kind == "reference":
        ref = oracle.reference()
        d1 = torch.zeros(a.b, hi - lo)
        d2 = torch.zeros(a.b, a.m)
        i1 = torch.zeros(a.b, hi - lo, dtype=torch.int32)
        i2 = torch.zeros(a.b, a.m, dtype=torch.int32)
        g1, g2 = torch.zeros_like(q), torch.zeros_like(x2)

        def step():
            ref.nnd_forward(q, x2, d1, d2, i1, i2)  # my_lib.cpp:28-60, both directions
            ref.nnd_backward(q, x2, g1, g2, gd1, gd2, i1, i2)  # my_lib.cpp:64-133
    else:
        qn, tn = q.numpy(), x2.numpy()

        def step():
            d1, d2, i1, i2 = oracle.forward(qn, tn)
            oracle.backward(qn, tn, gd1.numpy(), gd2.numpy(), i1, i2)

    while time.time() < a.start_at:
        time.sleep(0.005)
    times = []
    for s in range(a.warmup + a.steps):
        t0 = time.perf_counter()
        step()
        if s >= a.warmup:
            times.append(time.perf_counter() - t0)
    print(json.dumps({"kind": kind, "step_seconds": times, "rows": hi - lo}))


if __name__ == "__main__":
    main()
