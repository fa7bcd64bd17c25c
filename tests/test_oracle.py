"""CPU tests: the oracle (oracle/nnd_oracle.c) against the reference-generated golden vectors,
against the reference's own compiled my_lib.cpp when oracle/_ref is present, and against a dense
torch formulation.  No GPU, no product code under test here -- this pins the checker."""
import numpy as np
import pytest
import torch

import oracle
from _util import DIGEST_CASES, DIGESTS, NPZ_CASES, assert_forward_exact, bits_equal, load_case, sha
from pointcloudregistration_b200 import synthetic


@pytest.mark.parametrize("name", NPZ_CASES)
def test_port_matches_golden_vectors(name):
    _, z = load_case(name)
    got = oracle.forward(z["xyz1"], z["xyz2"])
    assert_forward_exact(got, (z["dist1"], z["dist2"], z["idx1"], z["idx2"]), what=name + ": ")
    g1, g2 = oracle.backward(z["xyz1"], z["xyz2"], z["graddist1"], z["graddist2"], z["idx1"], z["idx2"])
    assert bits_equal(g1, z["grad1"]) and bits_equal(g2, z["grad2"])  # same sequential order -> bit-exact


@pytest.mark.parametrize("name", DIGEST_CASES)
def test_port_matches_golden_digests(name):
    meta, z = load_case(name)
    d1, d2, i1, i2 = oracle.forward(z["xyz1"], z["xyz2"], threads=4)
    g1, g2 = oracle.backward(z["xyz1"], z["xyz2"], z["graddist1"], z["graddist2"], i1, i2)
    got = dict(dist1=d1, dist2=d2, idx1=i1, idx2=i2, grad1=g1, grad2=g2)
    for k, v in got.items():
        assert sha(v) == meta["sha256"][k], f"{name}: {k} digest differs from the reference run"


@pytest.mark.skipif(not oracle.have_reference(), reason="oracle/_ref not built (needs /root/reference)")
@pytest.mark.parametrize("kind", ["uniform", "lattice", "duplicates", "crosscopy", "clustered", "large", "tiny"])
def test_port_matches_compiled_reference(kind):
    x1, x2 = synthetic.make_pair(kind, 2, 700, 450, 31)
    gd1, gd2 = synthetic.grad_pair(2, 700, 450, 31)
    want = oracle.reference_forward(x1.numpy(), x2.numpy())
    got = oracle.forward(x1.numpy(), x2.numpy())
    assert_forward_exact(got, want, what=kind + ": ")
    rg = oracle.reference_backward(x1.numpy(), x2.numpy(), gd1.numpy(), gd2.numpy(), want[2], want[3])
    pg = oracle.backward(x1.numpy(), x2.numpy(), gd1.numpy(), gd2.numpy(), got[2], got[3])
    assert bits_equal(pg[0], rg[0]) and bits_equal(pg[1], rg[1])


@pytest.mark.skipif(not oracle.have_reference(), reason="oracle/_ref not built (needs /root/reference)")
def test_port_matches_compiled_reference_on_nonfinite_inputs():
    """my_lib.cpp:16 accepts k == 0 unconditionally: NaN sticks only when it comes first; +inf competes normally."""
    x1 = synthetic.uniform(2, 300, 41).numpy()
    x2 = synthetic.uniform(2, 200, 42).numpy()
    x1[0, 3] = np.nan
    x1[1, 0, 1] = np.nan
    x1[1, 17] = 3e19
    x2[0, 0] = np.inf
    x2[0, 9, 1] = np.nan
    x2[1, 100] = -np.inf
    want = oracle.reference_forward(x1, x2)
    got = oracle.forward(x1, x2)
    for g, w in zip(got[2:], want[2:]):
        assert np.array_equal(g, w)
    for g, w in zip(got[:2], want[:2]):
        assert np.array_equal(np.isnan(g), np.isnan(w)) and g[~np.isnan(g)].tobytes() == w[~np.isnan(w)].tobytes()
    assert np.isnan(got[0]).any() and np.isinf(got[0]).any()


def _dense(x1, x2):
    # (d0 + d1) + d2 with every op rounded to fp32, first-occurrence argmin: my_lib.cpp:12-19
    diff = x2[:, None, :, :] - x1[:, :, None, :]
    sq = diff * diff
    d = (sq[..., 0] + sq[..., 1]) + sq[..., 2]
    return d


def test_port_matches_dense_torch():
    x1, x2 = synthetic.make_pair("lattice", 2, 300, 211, 32)
    d = _dense(x1, x2)
    d1, i1 = d.min(dim=2)
    d2, i2 = d.min(dim=1)
    got = oracle.forward(x1.numpy(), x2.numpy())
    assert bits_equal(got[0], d1.numpy()) and bits_equal(got[1], d2.numpy())
    # torch.min does not promise first-occurrence on ties; check idx through its defining property instead
    dn = d.numpy()
    for b in range(2):
        for j in range(300):
            k = got[2][b, j]
            assert dn[b, j, k] == got[0][b, j] and not (dn[b, j, :k] <= got[0][b, j]).any()
        for k in range(211):
            j = got[3][b, k]
            assert dn[b, j, k] == got[1][b, k] and not (dn[b, :j, k] <= got[1][b, k]).any()


def test_backward_matches_autograd_of_dense_formulation():
    x1, x2 = synthetic.make_pair("uniform", 2, 120, 90, 33)
    gd1, gd2 = synthetic.grad_pair(2, 120, 90, 33)
    a = x1.clone().double().requires_grad_(True)
    b = x2.clone().double().requires_grad_(True)
    d = ((b[:, None] - a[:, :, None]) ** 2).sum(-1)
    loss = (d.min(2).values * gd1.double()).sum() + (d.min(1).values * gd2.double()).sum()
    loss.backward()
    _, _, i1, i2 = oracle.forward(x1.numpy(), x2.numpy())
    g1, g2 = oracle.backward(x1.numpy(), x2.numpy(), gd1.numpy(), gd2.numpy(), i1, i2)
    np.testing.assert_allclose(g1, a.grad.numpy(), rtol=1e-4, atol=1e-6)
    np.testing.assert_allclose(g2, b.grad.numpy(), rtol=1e-4, atol=1e-6)


def test_multithreaded_driver_is_identical():
    x1, x2 = synthetic.make_pair("duplicates", 3, 1001, 777, 34)
    assert_forward_exact(oracle.forward(x1.numpy(), x2.numpy(), threads=5), oracle.forward(x1.numpy(), x2.numpy()))


def test_empty_target_gives_zero_like_reference():
    # my_lib.cpp:9-10,21-22: with m == 0 the inner loop never runs -> best = 0, besti = 0
    x1 = synthetic.uniform(1, 5, 1).numpy()
    x2 = np.zeros((1, 0, 3), np.float32)
    d1, d2, i1, i2 = oracle.forward(x1, x2)
    assert (d1 == 0).all() and (i1 == 0).all() and d2.shape == (1, 0) and i2.shape == (1, 0)


def test_ties_resolve_to_lowest_index():
    q = np.zeros((1, 1, 3), np.float32)
    t = np.array([[[2, 0, 0], [1, 0, 0], [0, 1, 0], [0, 0, 1], [1, 0, 0]]], np.float32)
    d, i = oracle.query(q, t)
    assert d[0, 0] == 1.0 and i[0, 0] == 1


def test_no_fma_contraction_in_oracle_build():
    # (x*x + y*y) + z*z with separately rounded products differs from the fused form for these values
    x = np.float32(1.0000001)
    q = np.zeros((1, 1, 3), np.float32)
    t = np.array([[[x, x, x]]], np.float32)
    d, _ = oracle.query(q, t)
    xx = np.float32(x * x)
    assert d[0, 0] == np.float32(np.float32(xx + xx) + xx)


def test_golden_digest_table_is_complete():
    for name, meta in DIGESTS.items():
        assert set(meta["sha256"]) == {"dist1", "dist2", "idx1", "idx2", "grad1", "grad2"}, name


@pytest.mark.skipif(not oracle.have_reference(), reason="oracle/_ref not built (needs /root/reference)")
def test_port_matches_compiled_reference_on_random_shapes():
    """Property test (hypothesis): any small shape, coordinates drawn from a coarse grid so that exact ties and
    duplicates are everywhere -- restatement and compiled reference agree on every bit, forward and backward."""
    from hypothesis import given, settings
    from hypothesis import strategies as st

    @settings(max_examples=40, deadline=None, derandomize=True)
    @given(b=st.integers(1, 3), n=st.integers(1, 90), m=st.integers(1, 90), cells=st.sampled_from([2, 5, 64, 4096]),
           seed=st.integers(0, 2**31 - 1))
    def check(b, n, m, cells, seed):
        rng = np.random.default_rng(seed)
        x1 = (rng.integers(0, cells, size=(b, n, 3)) / np.float32(cells)).astype(np.float32)
        x2 = (rng.integers(0, cells, size=(b, m, 3)) / np.float32(cells)).astype(np.float32)
        gd1 = rng.standard_normal((b, n)).astype(np.float32)
        gd2 = rng.standard_normal((b, m)).astype(np.float32)
        want = oracle.reference_forward(x1, x2)
        got = oracle.forward(x1, x2)
        assert_forward_exact(got, want)
        rg = oracle.reference_backward(x1, x2, gd1, gd2, want[2], want[3])
        pg = oracle.backward(x1, x2, gd1, gd2, got[2], got[3])
        assert bits_equal(pg[0], rg[0]) and bits_equal(pg[1], rg[1])

    check()
