"""tests/golden/make_golden_full.py -- full-size golden digests (BASELINE.json configs 2-5), dev container only.

Runs the REFERENCE's own CPU nnd (my_lib.cpp compiled unmodified into oracle/_ref by oracle/build_ref.py) over the
WHOLE clouds of the full-size configs and stores SHA-256 digests of every output tensor in digests_full.json, so
the GPU tests (and bench.py's parity block) compare complete tensors, not row samples.

The reference is single-threaded (2e12 pairs at C4 = ~1.5 h on one core), so its own entry point nnd_forward is
called on disjoint slices of cloud-1 rows against the full cloud 2, one slice per worker process:
  * dist1/idx1 of a slice are final (nnsearch is a pure per-row function, my_lib.cpp:3-25);
  * dist2/idx2 of a slice call are the minima over that slice's rows only; slices are merged in ascending row order
    with a strict `<`, which is exactly the order and the comparison of the k loop (my_lib.cpp:11-20), so the merged
    result is what one call on the whole clouds returns.  `--selfcheck` proves that on cases small enough for a
    single whole-cloud call of the reference.
nnd_backward (my_lib.cpp:64-133) runs once on the full tensors (it is O(N)).

    python tests/golden/make_golden_full.py [--workers 7] [--only c5_lattice ...] [--selfcheck]
"""
from __future__ import annotations

import argparse
import hashlib
import json
import multiprocessing as mp
import sys
import time
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parents[2]
sys.path.insert(0, str(ROOT))

HERE = Path(__file__).resolve().parent
OUT = HERE / "digests_full.json"

# name -> (kind, b, n, m, config number = seeds)
CASES = {
    "c1_uniform": ("uniform", 1, 2048, 2048, 1),
    "dip_uniform": ("uniform", 256, 256, 256, 6),  # the shape DIP really trains on (dip/train.py:84)
    "c2_uniform": ("uniform", 32, 16384, 16384, 2),
    "c3_uniform": ("uniform", 1, 100_000, 100_000, 3),
    "c5_lattice": ("lattice", 1, 262_144, 262_144, 5),
    "c5_duplicates": ("duplicates", 1, 262_144, 262_144, 5),
    "c5_identical": ("identical", 1, 262_144, 262_144, 5),
    "c5_crosscopy": ("crosscopy", 1, 262_144, 262_144, 5),
    "c5_clustered": ("clustered", 1, 262_144, 262_144, 5),
    "c4_uniform": ("uniform", 1, 1_000_000, 1_000_000, 4),
    "c4_clustered": ("clustered", 1, 1_000_000, 1_000_000, 4),
}

_clouds = {}


def _pair(kind, b, n, m, cfg):
    from pointcloudregistration_b200 import synthetic

    key = (kind, b, n, m, cfg)
    if key not in _clouds:
        _clouds.clear()
        _clouds[key] = synthetic.make_pair(kind, b, n, m, cfg)
    return _clouds[key]


def _slice_job(job):
    """One call of the reference's nnd_forward on rows [lo, hi) of cloud 1 against the full cloud 2."""
    import torch

    torch.set_num_threads(1)
    import oracle

    kind, b, n, m, cfg, lo, hi = job
    x1, x2 = _pair(kind, b, n, m, cfg)
    q = x1[:, lo:hi].contiguous()
    d1 = torch.zeros(b, hi - lo)
    d2 = torch.zeros(b, m)
    i1 = torch.zeros(b, hi - lo, dtype=torch.int32)
    i2 = torch.zeros(b, m, dtype=torch.int32)
    oracle.reference().nnd_forward(q, x2, d1, d2, i1, i2)
    return lo, hi, d1.numpy(), i1.numpy(), d2.numpy(), i2.numpy()


def sliced_reference_forward(pool, kind, b, n, m, cfg, rows_per_job):
    jobs = [(kind, b, n, m, cfg, lo, min(n, lo + rows_per_job)) for lo in range(0, n, rows_per_job)]
    d1 = np.empty((b, n), np.float32)
    i1 = np.empty((b, n), np.int32)
    parts = {}
    for lo, hi, pd1, pi1, pd2, pi2 in pool.imap_unordered(_slice_job, jobs):
        d1[:, lo:hi] = pd1
        i1[:, lo:hi] = pi1
        parts[lo] = (pd2, pi2 + np.int32(lo))
    d2 = i2 = None
    for lo in sorted(parts):  # ascending rows, strict <: the k loop's own order (my_lib.cpp:11-20)
        pd2, pi2 = parts[lo]
        if d2 is None:
            d2, i2 = pd2.copy(), pi2.copy()
        else:
            better = pd2 < d2
            d2[better] = pd2[better]
            i2[better] = pi2[better]
    return d1, d2, i1, i2


def sha(a) -> str:
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


def run_case(pool, name, rows_per_job):
    import oracle
    from pointcloudregistration_b200 import synthetic

    kind, b, n, m, cfg = CASES[name]
    t0 = time.time()
    x1, x2 = synthetic.make_pair(kind, b, n, m, cfg)
    d1, d2, i1, i2 = sliced_reference_forward(pool, kind, b, n, m, cfg, rows_per_job)
    gd1, gd2 = synthetic.grad_pair(b, n, m, cfg)
    g1, g2 = oracle.reference_backward(x1.numpy(), x2.numpy(), gd1.numpy(), gd2.numpy(), i1, i2)
    md1, md2 = synthetic.grad_pair(b, n, m, cfg, mean=True)  # what bench.py feeds (mean chamfer)
    mg1, mg2 = oracle.reference_backward(x1.numpy(), x2.numpy(), md1.numpy(), md2.numpy(), i1, i2)
    out = dict(dist1=d1, dist2=d2, idx1=i1, idx2=i2, grad1=g1, grad2=g2, grad1_mean=mg1, grad2_mean=mg2)
    inputs = dict(xyz1=x1.numpy(), xyz2=x2.numpy(), graddist1=gd1.numpy(), graddist2=gd2.numpy())
    rec = dict(kind=kind, b=b, n=n, m=m, config=cfg, sha256={k: sha(v) for k, v in out.items()},
               input_sha256={k: sha(v) for k, v in inputs.items()},
               # float64 sums: a cheap second handle when a digest mismatches
               sums=dict(dist1=float(d1.astype(np.float64).sum()), dist2=float(d2.astype(np.float64).sum()),
                         idx1=int(i1.astype(np.int64).sum()), idx2=int(i2.astype(np.int64).sum())),
               generator="reference my_lib.cpp compiled unmodified (oracle/_ref), nnd_forward on cloud-1 row slices of "
                         f"{rows_per_job} rows, dist2/idx2 merged in ascending row order with strict <; nnd_backward on "
                         "the whole tensors",
               seconds=round(time.time() - t0, 1))
    return rec


def selfcheck(pool):
    """The sliced + merged forward equals one whole-cloud call of the reference (ties, duplicates, several slices)."""
    import oracle
    from pointcloudregistration_b200 import synthetic

    for kind, b, n, m in (("lattice", 2, 5000, 4000), ("identical", 1, 6000, 6000), ("crosscopy", 1, 3000, 7000)):
        x1, x2 = synthetic.make_pair(kind, b, n, m, 77)
        want = oracle.reference_forward(x1.numpy(), x2.numpy())
        got = sliced_reference_forward(pool, kind, b, n, m, 77, 700)
        for w, g, nm in zip(want, got, ("dist1", "dist2", "idx1", "idx2")):
            assert w.tobytes() == g.tobytes(), f"selfcheck {kind}: {nm} differs"
    print("selfcheck ok: sliced + merged == whole-cloud reference call")


def main():
    import oracle

    ap = argparse.ArgumentParser()
    ap.add_argument("--workers", type=int, default=7)
    ap.add_argument("--only", nargs="*", default=None)
    ap.add_argument("--rows-per-job", type=int, default=4096)
    ap.add_argument("--selfcheck", action="store_true")
    a = ap.parse_args()
    assert oracle.have_reference(), "build oracle/_ref first: python oracle/build_ref.py"
    digests = json.loads(OUT.read_text()) if OUT.exists() else {}
    with mp.get_context("fork").Pool(a.workers) as pool:
        if a.selfcheck:
            selfcheck(pool)
        for name in (a.only if a.only is not None else CASES):
            rec = run_case(pool, name, a.rows_per_job)
            digests[name] = rec
            OUT.write_text(json.dumps(digests, indent=1, sort_keys=True) + "\n")  # after every case: hours of CPU
            print(f"{name}: {rec['seconds']} s", flush=True)


if __name__ == "__main__":
    main()
