"""tests/golden/make_golden.py -- regenerates tests/golden/*.npz and digests.json.

Runs the REFERENCE's own CPU nnd (my_lib.cpp compiled unmodified into oracle/_ref by
oracle/build_ref.py; needs /root/reference, i.e. the dev container) on seeded inputs and
stores (a) small cases in full (inputs are regenerated from the seed, outputs stored) and
(b) SHA-256 digests of the output bytes for medium cases.  The reference ships no golden
vectors for this path (dip/torch-nndistance/test.py only prints), so these files ARE the
pin: tests/test_oracle.py checks the C restatement against them everywhere, and the GPU
parity tests check the CUDA path against them on the B200 box.

    python tests/golden/make_golden.py
"""
from __future__ import annotations

import hashlib
import json
import sys
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parents[2]
sys.path.insert(0, str(ROOT))

import oracle  # noqa: E402
from pointcloudregistration_b200 import synthetic  # noqa: E402

HERE = Path(__file__).resolve().parent

# (name, kind, b, n, m, config-number-for-seeds)
SMALL = [
    ("uniform_b2_n257_m130", "uniform", 2, 257, 130, 11),
    ("lattice_b1_n300_m300", "lattice", 1, 300, 300, 12),
    ("identical_b1_n256", "identical", 1, 256, 256, 13),
    ("clustered_b3_n64_m500", "clustered", 3, 64, 500, 14),
    ("single_target_b2_n33_m1", "uniform", 2, 33, 1, 15),
    ("large_b1_n200_m333", "large", 1, 200, 333, 16),
    ("tiny_b1_n128_m128", "tiny", 1, 128, 128, 17),
]
DIGEST = [
    ("c1_uniform_b1_n2048_m2048", "uniform", 1, 2048, 2048, 1),
    ("smoke_shape_b16_n2048_m1024", "uniform", 16, 2048, 1024, 21),  # dip/torch-nndistance/test.py:8-9
    ("lattice_b1_n4096_m4096", "lattice", 1, 4096, 4096, 22),
    ("duplicates_b2_n3000_m2500", "duplicates", 2, 3000, 2500, 23),
    ("crosscopy_b1_n5000_m4097", "crosscopy", 1, 5000, 4097, 24),
    ("clustered_b1_n8192_m8192", "clustered", 1, 8192, 8192, 25),
    ("dip_patch_b64_n256_m256", "uniform", 64, 256, 256, 26),  # real DIP shape, dip/train.py:84
]


def run_reference(kind, b, n, m, cfg):
    x1, x2 = synthetic.make_pair(kind, b, n, m, cfg)
    gd1, gd2 = synthetic.grad_pair(b, n, m, cfg)
    d1, d2, i1, i2 = oracle.reference_forward(x1.numpy(), x2.numpy())
    g1, g2 = oracle.reference_backward(x1.numpy(), x2.numpy(), gd1.numpy(), gd2.numpy(), i1, i2)
    out = dict(dist1=d1, dist2=d2, idx1=i1, idx2=i2, grad1=g1, grad2=g2)
    inputs = dict(xyz1=x1.numpy(), xyz2=x2.numpy(), graddist1=gd1.numpy(), graddist2=gd2.numpy())
    return out, inputs


def digest(out) -> dict:
    return {k: hashlib.sha256(np.ascontiguousarray(v).tobytes()).hexdigest() for k, v in out.items()}


def main():
    assert oracle.have_reference(), "build oracle/_ref first: python oracle/build_ref.py"
    digests = {}
    for name, kind, b, n, m, cfg in SMALL:
        out, inputs = run_reference(kind, b, n, m, cfg)
        np.savez_compressed(HERE / f"{name}.npz", **out, **inputs)  # inputs stored too: no RNG dependence
        digests[name] = dict(kind=kind, b=b, n=n, m=m, config=cfg, stored="npz", sha256=digest(out))
    for name, kind, b, n, m, cfg in DIGEST:
        out, inputs = run_reference(kind, b, n, m, cfg)
        digests[name] = dict(kind=kind, b=b, n=n, m=m, config=cfg, stored="digest", sha256=digest(out),
                             input_sha256=digest(inputs))
    (HERE / "digests.json").write_text(json.dumps(digests, indent=1, sort_keys=True) + "\n")
    print(f"wrote {len(SMALL)} npz files and {len(digests)} digests to {HERE}")


if __name__ == "__main__":
    main()
