"""Shared helpers for the parity tests."""
from __future__ import annotations

import hashlib
import json
from pathlib import Path

import numpy as np

GOLDEN = Path(__file__).resolve().parent / "golden"
DIGESTS = json.loads((GOLDEN / "digests.json").read_text())
NPZ_CASES = sorted(k for k, v in DIGESTS.items() if v["stored"] == "npz")
DIGEST_CASES = sorted(k for k, v in DIGESTS.items() if v["stored"] == "digest")
# full-size configs (C2-C5): digests of the WHOLE output tensors of the compiled reference (make_golden_full.py)
FULL_DIGESTS = json.loads((GOLDEN / "digests_full.json").read_text())

# gradient tolerance north_star states: 1e-5 relative (summation order of the atomics differs)
GRAD_RTOL = 1e-5


def sha(a) -> str:
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


def bits_equal(a, b) -> bool:
    a = np.ascontiguousarray(a)
    b = np.ascontiguousarray(b)
    return a.shape == b.shape and a.dtype == b.dtype and a.tobytes() == b.tobytes()


def assert_forward_exact(got, want, what=""):
    """idx bit-exact, dist bit-exact in fp32."""
    names = ("dist1", "dist2", "idx1", "idx2")
    for name, g, w in zip(names, got, want):
        g = np.asarray(g)
        w = np.asarray(w)
        assert g.shape == w.shape, f"{what}{name}: shape {g.shape} vs {w.shape}"
        if not bits_equal(g, w):
            bad = np.flatnonzero(g.reshape(-1).view(np.int32) != w.reshape(-1).view(np.int32))
            i = int(bad[0])
            raise AssertionError(f"{what}{name}: {bad.size} of {g.size} entries differ; first at flat index {i}: "
                                 f"got {g.reshape(-1)[i]!r} want {w.reshape(-1)[i]!r}")


def assert_grad_close(got, want, what="", rtol=GRAD_RTOL):
    """|got - want| <= rtol * max|want| per tensor, plus an element-wise rtol/atol check.

    The oracle sums scatter contributions sequentially (my_lib.cpp:96-129); the GPU sums them with
    atomics in arbitrary order, so only a relative bound holds.  The scale is the largest gradient
    magnitude of the tensor: an entry that cancels to ~0 cannot be held to 1e-5 of itself."""
    for name, g, w in zip(("gradxyz1", "gradxyz2"), got, want):
        g = np.asarray(g, dtype=np.float64)
        w = np.asarray(w, dtype=np.float64)
        assert g.shape == w.shape, f"{what}{name}: shape {g.shape} vs {w.shape}"
        scale = float(np.max(np.abs(w))) if w.size else 0.0
        err = float(np.max(np.abs(g - w))) if w.size else 0.0
        assert err <= rtol * max(scale, 1e-30), f"{what}{name}: max abs err {err:.3e} > {rtol} * {scale:.3e}"


def load_case(name):
    """(meta, inputs/outputs dict).  npz cases carry everything; digest cases regenerate inputs from seeds."""
    from pointcloudregistration_b200 import synthetic

    meta = DIGESTS[name]
    if meta["stored"] == "npz":
        z = np.load(GOLDEN / f"{name}.npz")
        return meta, {k: z[k] for k in z.files}
    x1, x2 = synthetic.make_pair(meta["kind"], meta["b"], meta["n"], meta["m"], meta["config"])
    gd1, gd2 = synthetic.grad_pair(meta["b"], meta["n"], meta["m"], meta["config"])
    data = dict(xyz1=x1.numpy(), xyz2=x2.numpy(), graddist1=gd1.numpy(), graddist2=gd2.numpy())
    for k, v in data.items():  # the seeded generator must still produce the inputs the digests were made from
        assert sha(v) == meta["input_sha256"][k], f"{name}: regenerated input {k} differs from the golden run"
    return meta, data
