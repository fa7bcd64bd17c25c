"""GPU parity tests (run with -m gpu on a B200): the CUDA path, called through the C ABI
(ctypes -> libnnd_b200.so), against the CPU oracle and the reference-generated golden vectors.

Bar (BASELINE.json north_star): idx bit-exact, dist bit-exact in fp32, gradients within 1e-5 relative."""
import numpy as np
import pytest
import torch

import oracle
from _util import (DIGEST_CASES, FULL_DIGESTS, GRAD_RTOL, NPZ_CASES, assert_forward_exact, assert_grad_close, load_case,
                   sha)
from pointcloudregistration_b200 import functional as F
from pointcloudregistration_b200 import synthetic

pytestmark = pytest.mark.gpu


def gpu_forward(x1, x2, device):
    out = F.nnd_forward_raw(torch.as_tensor(x1).to(device), torch.as_tensor(x2).to(device))
    return tuple(t.cpu().numpy() for t in out)


def gpu_backward(x1, x2, gd1, gd2, i1, i2, device, exact=False):
    t = [torch.as_tensor(np.asarray(a)).to(device) for a in (x1, x2, gd1, gd2, i1, i2)]
    g1, g2 = F.nnd_backward_raw(*t, exact=exact)
    return g1.cpu().numpy(), g2.cpu().numpy()


def assert_grad_bits(got, want, what=""):
    """The deterministic backward: every gradient bit equal to the reference CPU path's."""
    for name, g, w in zip(("gradxyz1", "gradxyz2"), got, want):
        g, w = np.ascontiguousarray(g), np.ascontiguousarray(w, dtype=np.float32)
        assert g.shape == w.shape, f"{what}{name}: shape {g.shape} vs {w.shape}"
        bad = np.flatnonzero(g.reshape(-1).view(np.int32) != w.reshape(-1).view(np.int32))
        assert bad.size == 0, (f"{what}{name}: {bad.size} of {g.size} entries differ in bits; first at {int(bad[0])}: "
                               f"got {g.reshape(-1)[bad[0]]!r} want {w.reshape(-1)[bad[0]]!r}")


# ---------------------------------------------------------------- golden vectors (reference outputs)
@pytest.mark.parametrize("name", NPZ_CASES)
def test_golden_vectors(name, cuda_lib, device):
    _, z = load_case(name)
    got = gpu_forward(z["xyz1"], z["xyz2"], device)
    assert_forward_exact(got, (z["dist1"], z["dist2"], z["idx1"], z["idx2"]), what=name + ": ")
    g = gpu_backward(z["xyz1"], z["xyz2"], z["graddist1"], z["graddist2"], z["idx1"], z["idx2"], device)
    assert_grad_close(g, (z["grad1"], z["grad2"]), what=name + ": ")
    g = gpu_backward(z["xyz1"], z["xyz2"], z["graddist1"], z["graddist2"], z["idx1"], z["idx2"], device, exact=True)
    assert_grad_bits(g, (z["grad1"], z["grad2"]), what=name + " (deterministic backward): ")


@pytest.mark.parametrize("name", DIGEST_CASES)
def test_golden_digests(name, cuda_lib, device):
    meta, z = load_case(name)
    d1, d2, i1, i2 = gpu_forward(z["xyz1"], z["xyz2"], device)
    for k, v in dict(dist1=d1, dist2=d2, idx1=i1, idx2=i2).items():
        assert sha(v) == meta["sha256"][k], f"{name}: {k} differs from the reference run"


# ---------------------------------------------------------------- shape / data matrix vs the oracle
SHAPES = [
    (1, 1, 1), (1, 1, 7), (1, 7, 1), (2, 31, 33), (1, 32, 32), (1, 33, 31), (3, 511, 513), (1, 512, 512),
    (1, 2048, 2048), (2, 2049, 1000), (1, 100, 5000), (1, 5000, 100), (5, 700, 300), (1, 4097, 6001),
    (40, 64, 70), (1, 20000, 1500),
]


@pytest.fixture(params=["auto", "general"])
def fwd_path(request, cuda_lib):
    """auto: clouds of <= 4096 points take the single-launch kernels (nnd_small.cuh).  general: that path switched
    off, so the same shapes exercise prepare + nnd_scan_kernel + finalize and the memset + atomic backward."""
    before = cuda_lib.nnd_set_small_max_points(0 if request.param == "general" else -1)
    yield request.param
    cuda_lib.nnd_set_small_max_points(before)


@pytest.mark.parametrize("b,n,m", SHAPES)
def test_shapes_uniform(b, n, m, fwd_path, cuda_lib, device):
    x1, x2 = synthetic.make_pair("uniform", b, n, m, 40)
    want = oracle.forward(x1.numpy(), x2.numpy(), threads=8)
    small = fwd_path == "auto" and (max(n, m) <= 4096 or (b == 1 and max(n, m) <= 10240))
    assert cuda_lib.nnd_forward_kind(b, n, m) == (0 if small else 1)
    assert_forward_exact(gpu_forward(x1, x2, device), want, what=f"b{b} n{n} m{m}: ")
    assert cuda_lib.nnd_last_launch_count() == (1 if small else 3)


@pytest.mark.parametrize("kind", ["clustered", "lattice", "duplicates", "identical", "crosscopy", "large", "tiny"])
@pytest.mark.parametrize("b,n,m", [(2, 1500, 1500), (1, 6000, 6000)])
def test_data_kinds(kind, b, n, m, fwd_path, cuda_lib, device):
    x1, x2 = synthetic.make_pair(kind, b, n, m, 41)
    want = oracle.forward(x1.numpy(), x2.numpy(), threads=8)
    got = gpu_forward(x1, x2, device)
    assert_forward_exact(got, want, what=f"{kind}: ")
    gd1, gd2 = synthetic.grad_pair(b, n, m, 41)
    wg = oracle.backward(x1.numpy(), x2.numpy(), gd1.numpy(), gd2.numpy(), want[2], want[3])
    assert_grad_close(gpu_backward(x1, x2, gd1, gd2, got[2], got[3], device), wg, what=f"{kind}: ")
    assert_grad_bits(gpu_backward(x1, x2, gd1, gd2, got[2], got[3], device, exact=True), wg, what=f"{kind} (exact): ")


# ---------------------------------------------------------------- single-launch small-cloud kernels (nnd_small.cuh)
SMALL_CONFIGS = [(w, q, t) for w in (2, 4, 8, 16) for q in (1, 2, 4, 8) for t in (1, 2, 4, 8, 16) if w % t == 0]


@pytest.mark.parametrize("warps,qpt,tsplit", SMALL_CONFIGS)
def test_small_kernel_every_configuration(warps, qpt, tsplit, cuda_lib, device):
    """Every (warps per CTA, queries per lane, target split) of nnd_small_kernel gives the oracle's bits: ragged
    shapes, a single target, exact ties (lattice: lowest index must win across target splits too), duplicates."""
    assert cuda_lib.nnd_set_small_config(warps, qpt, tsplit) == 0
    try:
        for kind, b, n, m in (("uniform", 3, 257, 130), ("lattice", 1, 2048, 2048), ("crosscopy", 2, 3000, 2900),
                              ("uniform", 1, 33, 4096), ("uniform", 5, 1, 7), ("identical", 2, 700, 700),
                              ("uniform", 2, 40, 1), ("duplicates", 64, 256, 256)):
            x1, x2 = synthetic.make_pair(kind, b, n, m, 50)
            want = oracle.forward(x1.numpy(), x2.numpy(), threads=8)
            assert_forward_exact(gpu_forward(x1, x2, device), want, what=f"cfg {warps},{qpt},{tsplit} {kind} b{b} n{n} m{m}: ")
            assert cuda_lib.nnd_last_launch_count() == 1
            q, t = x1.to(device), x2.to(device)
            d, i = F.nnd_query_raw(q, t)  # the one-direction entry point shares the kernel
            assert d.cpu().numpy().tobytes() == want[0].tobytes() and i.cpu().numpy().tobytes() == want[2].tobytes()
    finally:
        cuda_lib.nnd_set_small_config(0, 0, 0)


def test_small_kernel_up_to_its_hard_maximum(cuda_lib, device):
    """Limit raised to 16384 points: 3 x 16384 x 4 B = 192 KB of shared memory per CTA (opt-in above 48 KB)."""
    before = cuda_lib.nnd_set_small_max_points(16384)
    try:
        for kind, b, n, m in (("uniform", 1, 9000, 12000), ("lattice", 2, 16384, 5000), ("clustered", 1, 16384, 16384)):
            x1, x2 = synthetic.make_pair(kind, b, n, m, 51)
            assert cuda_lib.nnd_forward_kind(b, n, m) == 0
            want = oracle.forward(x1.numpy(), x2.numpy(), threads=16)
            got = gpu_forward(x1, x2, device)
            assert_forward_exact(got, want, what=f"{kind} b{b} n{n} m{m}: ")
            assert cuda_lib.nnd_last_launch_count() == 1
            if True:
                gd1, gd2 = synthetic.grad_pair(b, n, m, 51)
                wg = oracle.backward(x1.numpy(), x2.numpy(), gd1.numpy(), gd2.numpy(), want[2], want[3])
                assert_grad_close(gpu_backward(x1, x2, gd1, gd2, got[2], got[3], device), wg)
                assert cuda_lib.nnd_last_launch_count() == 1
    finally:
        cuda_lib.nnd_set_small_max_points(before)


@pytest.mark.parametrize("kind,b,n,m", [("uniform", 1, 1, 1), ("uniform", 256, 256, 256), ("clustered", 4, 3000, 2900),
                                        ("identical", 2, 2048, 2048), ("uniform", 1, 4096, 4096),
                                        ("crosscopy", 7, 100, 1900)])
def test_small_backward_is_one_launch_and_matches_oracle(kind, b, n, m, cuda_lib, device):
    x1, x2 = synthetic.make_pair(kind, b, n, m, 52)
    gd1, gd2 = synthetic.grad_pair(b, n, m, 52)
    _, _, i1, i2 = oracle.forward(x1.numpy(), x2.numpy(), threads=8)
    want = oracle.backward(x1.numpy(), x2.numpy(), gd1.numpy(), gd2.numpy(), i1, i2)
    t = [torch.as_tensor(np.asarray(a)).to(device) for a in (x1, x2, gd1, gd2, i1, i2)]
    import ctypes
    g1 = torch.full((b, n, 3), float("nan"), device=device)  # the call overwrites (my_lib.cpp:91-94)
    g2 = torch.full((b, m, 3), 7.0, device=device)
    rc = cuda_lib.nnd_backward_cuda_f32(*(a.data_ptr() for a in t), b, n, m, g1.data_ptr(), g2.data_ptr(),
                                        ctypes.c_void_p(torch.cuda.current_stream().cuda_stream))
    assert rc == 0 and cuda_lib.nnd_last_launch_count() == 1
    assert_grad_close((g1.cpu().numpy(), g2.cpu().numpy()), want, what=f"{kind} b{b} n{n} m{m}: ")


def test_small_backward_heavy_collisions(cuda_lib, device):
    """All 4000 rows of cloud 1 choose one point: 4000 shared-memory atomics on the same three addresses."""
    n, m = 4000, 64
    x1 = synthetic.uniform(1, n, 44) * 0.01
    x2 = synthetic.uniform(1, m, 45) + 10.0
    x2[0, 17] = 0.005
    gd1, gd2 = synthetic.grad_pair(1, n, m, 44)
    _, _, i1, i2 = oracle.forward(x1.numpy(), x2.numpy(), threads=8)
    assert (i1 == 17).all()
    want = oracle.backward(x1.numpy(), x2.numpy(), gd1.numpy(), gd2.numpy(), i1, i2)
    got = gpu_backward(x1, x2, gd1, gd2, i1, i2, device)
    assert cuda_lib.nnd_last_launch_count() == 1
    scale = np.abs(2 * gd1.numpy()[0, :, None] * (x1.numpy()[0] - x2.numpy()[0, 17])).sum(0).max()
    assert np.abs(got[1] - want[1]).max() <= GRAD_RTOL * scale
    assert_grad_close((got[0],), (want[0],))


def _ragged_reference(x1, x2, l1, l2, gd1=None, gd2=None):
    """Per-cloud oracle calls on the valid prefixes; zeros elsewhere (what one ragged launch must produce)."""
    b, n, m = x1.shape[0], x1.shape[1], x2.shape[1]
    d1, d2 = np.zeros((b, n), np.float32), np.zeros((b, m), np.float32)
    i1, i2 = np.zeros((b, n), np.int32), np.zeros((b, m), np.int32)
    g1, g2 = np.zeros((b, n, 3), np.float32), np.zeros((b, m, 3), np.float32)
    for k in range(b):
        a, c = int(l1[k]), int(l2[k])
        if a == 0 or c == 0:
            continue
        e1, e2, j1, j2 = oracle.forward(x1[k:k + 1, :a], x2[k:k + 1, :c])
        d1[k, :a], d2[k, :c], i1[k, :a], i2[k, :c] = e1[0], e2[0], j1[0], j2[0]
        if gd1 is not None:
            h1, h2 = oracle.backward(x1[k:k + 1, :a], x2[k:k + 1, :c], gd1[k:k + 1, :a], gd2[k:k + 1, :c], j1, j2)
            g1[k, :a], g2[k, :c] = h1[0], h2[0]
    return (d1, d2, i1, i2), (g1, g2)


@pytest.mark.parametrize("cfg", [(0, 0, 0), (8, 2, 4), (16, 1, 16)])
def test_ragged_batch_in_one_launch(cfg, cuda_lib, device):
    """Per-cloud lengths handed to the kernel (loss.py:143-144 passes lengths1/lengths2 to one knn_points call): the
    whole ragged batch is one forward launch and one backward launch; padded rows come back as zeros."""
    b, n, m = 6, 700, 900
    x1, x2 = synthetic.make_pair("crosscopy", b, n, m, 53)
    l1 = torch.tensor([700, 13, 0, 699, 1, 350])
    l2 = torch.tensor([900, 900, 5, 0, 1, 17])
    gd1, gd2 = synthetic.grad_pair(b, n, m, 53)
    want, wantg = _ragged_reference(x1.numpy(), x2.numpy(), l1, l2, gd1.numpy(), gd2.numpy())
    assert cuda_lib.nnd_set_small_config(*cfg) == 0
    try:
        a, c = x1.to(device), x2.to(device)
        got = F.nnd_forward_ragged_raw(a, c, l1.to(device), l2.to(device))
        assert cuda_lib.nnd_last_launch_count() == 1
        assert_forward_exact(tuple(t.cpu().numpy() for t in got), want, what="ragged: ")
        g = F.nnd_backward_ragged_raw(a, c, gd1.to(device), gd2.to(device), got[2], got[3], l1.to(device), l2.to(device))
        assert cuda_lib.nnd_last_launch_count() == 1
        assert_grad_close(tuple(t.cpu().numpy() for t in g), wantg, what="ragged backward: ")
        for k in range(b):  # rows past a length never receive gradient
            assert (g[0][k, int(l1[k]):] == 0).all() and (g[1][k, int(l2[k]):] == 0).all()
        # one direction, lengths on one side only, out-of-range lengths are clamped by the kernel
        d, i = F.nnd_query_ragged_raw(a, c, None, torch.tensor([900, 5000, 5, -3, 1, 17]).to(device))
        w2, _ = _ragged_reference(x1.numpy(), x2.numpy(), torch.full((b,), n), torch.tensor([900, 900, 5, 0, 1, 17]))
        assert d.cpu().numpy().tobytes() == w2[0].tobytes() and i.cpu().numpy().tobytes() == w2[2].tobytes()
    finally:
        cuda_lib.nnd_set_small_config(0, 0, 0)


# ---------------------------------------------------------------- deterministic backward (bit-identical to my_lib.cpp:91-130)
@pytest.mark.parametrize("kind,b,n,m", [
    ("uniform", 1, 1, 1), ("uniform", 3, 7, 5), ("uniform", 2, 2049, 1000), ("clustered", 1, 20000, 1500),
    ("identical", 1, 4096, 4096),      # every cloud-2 row is chosen by two rows: ordered pairs everywhere
    ("duplicates", 4, 3000, 2000), ("lattice", 1, 30000, 30000), ("clustered", 1, 262144, 200000),
])
def test_deterministic_backward_is_bit_identical_to_the_reference_cpu_path(kind, b, n, m, cuda_lib, device):
    x1, x2 = synthetic.make_pair(kind, b, n, m, 48)
    d1, d2, i1, i2 = gpu_forward(x1, x2, device)
    gd1, gd2 = synthetic.grad_pair(b, n, m, 48)
    want = oracle.backward(x1.numpy(), x2.numpy(), gd1.numpy(), gd2.numpy(), i1, i2)
    got = gpu_backward(x1, x2, gd1, gd2, i1, i2, device, exact=True)
    assert_grad_bits(got, want, what=f"{kind} b{b} n{n} m{m}: ")
    again = gpu_backward(x1, x2, gd1, gd2, i1, i2, device, exact=True)
    assert got[0].tobytes() == again[0].tobytes() and got[1].tobytes() == again[1].tobytes()


def test_deterministic_backward_one_neighbour_for_everybody_and_denormals(cuda_lib, device):
    """All 50 000 rows of cloud 1 choose the same point (one 50 000-entry segment, summed in order by one thread), and
    upstream gradients small enough that products and sums are denormal (the atomic path flushes those, this one
    must not)."""
    n, m = 50000, 40
    x1 = torch.rand(1, n, 3, generator=torch.Generator().manual_seed(5)) * 0.01
    x2 = torch.rand(1, m, 3, generator=torch.Generator().manual_seed(6)) + 5.0
    x2[0, 17] = 0.005
    d1, d2, i1, i2 = gpu_forward(x1, x2, device)
    assert (i1 == 17).all()
    for scale in (1.0, 1e-38):
        gd1 = torch.rand(1, n, generator=torch.Generator().manual_seed(7)) * scale
        gd2 = torch.rand(1, m, generator=torch.Generator().manual_seed(8)) * scale
        want = oracle.backward(x1.numpy(), x2.numpy(), gd1.numpy(), gd2.numpy(), i1, i2)
        assert_grad_bits(gpu_backward(x1, x2, gd1, gd2, i1, i2, device, exact=True), want, what=f"scale {scale}: ")


def test_deterministic_backward_through_autograd(cuda_lib, device):
    import torch_nndistance as NND
    from pointcloudregistration_b200 import set_exact_backward

    x1, x2 = synthetic.make_pair("clustered", 2, 5000, 4000, 49)
    a = x1.to(device).requires_grad_(True)
    c = x2.to(device).requires_grad_(True)
    set_exact_backward(True)
    try:
        d1, d2, i1, i2 = NND.nnd_with_idx(a, c)
        (d1.sum() * 0.25 + d2.sum() * 0.5).backward()
    finally:
        set_exact_backward(None)
    gd1 = np.full((2, 5000), 0.25, np.float32)
    gd2 = np.full((2, 4000), 0.5, np.float32)
    want = oracle.backward(x1.numpy(), x2.numpy(), gd1, gd2, i1.cpu().numpy(), i2.cpu().numpy())
    assert_grad_bits((a.grad.cpu().numpy(), c.grad.cpu().numpy()), want, what="autograd: ")


def test_adversarial_ties_across_chunks_and_tiles(cuda_lib, device):
    """Equal minima placed in different 32-target chunks, different 512-target tiles, different target
    splits, and at k = 0: the lowest index must win everywhere."""
    m = 70000
    t = np.full((1, m, 3), 5.0, np.float32)
    t[0, :, 0] += np.arange(m, dtype=np.float32) * 1e-3  # all far away, all different
    q = np.zeros((1, 6, 3), np.float32)
    spots = [[0, 31, 32], [5, 511, 512, 1024], [40000, 40001, 69999], [33, 65535], [69999], [1, 2, 3, 4]]
    for j, ks in enumerate(spots):
        q[0, j] = (j + 1) * 100.0
        for k in ks:
            t[0, k] = q[0, j] + np.array([0.25, 0, 0], np.float32)  # identical distance 0.0625
    d1, d2, i1, i2 = gpu_forward(q, t, device)
    assert i1[0].tolist() == [ks[0] for ks in spots]
    assert np.all(d1[0] == np.float32(0.0625))
    assert_forward_exact((d1, d2, i1, i2), oracle.forward(q, t, threads=8))


def test_rounding_is_unfused(cuda_lib, device):
    """One pair whose (x*x + y*y) + z*z differs between separately rounded and FMA-contracted evaluation."""
    rng = np.random.default_rng(5)
    q = rng.random((1, 4096, 3), dtype=np.float32)
    t = rng.random((1, 1, 3), dtype=np.float32) + 1.0
    d, _, _, _ = gpu_forward(q, t, device)
    diff = (t[0, 0][None, :] - q[0]).astype(np.float32)
    sq = (diff * diff).astype(np.float32)
    want = ((sq[:, 0] + sq[:, 1]).astype(np.float32) + sq[:, 2]).astype(np.float32)
    fused = np.float32(1) * (diff[:, 0].astype(np.float64) ** 2 + diff[:, 1].astype(np.float64) ** 2
                             + diff[:, 2].astype(np.float64) ** 2).astype(np.float32)
    assert (want != fused).any(), "test data does not discriminate"
    assert d[0].tobytes() == want.tobytes()


def test_nonfinite_inputs_do_not_hang(cuda_lib, device):
    x1 = synthetic.uniform(1, 600, 1).numpy()
    x2 = synthetic.uniform(1, 700, 2).numpy()
    x1[0, 3] = np.nan
    x2[0, 5] = np.inf
    x2[0, 9, 1] = np.nan
    d1, d2, i1, i2 = gpu_forward(x1, x2, device)
    assert d1.shape == (1, 600) and i2.shape == (1, 700)
    assert ((i1 >= 0) & (i1 < 700)).all() and ((i2 >= 0) & (i2 < 600)).all()
    # rows that never see a non-finite value still match the oracle exactly
    ok1 = np.ones(600, bool)
    ok1[3] = False
    w1, _ = oracle.query(np.delete(x1, 3, axis=1)[:, :, :], np.delete(x2, [5, 9], axis=1))
    g1, _ = oracle.query(x1, np.delete(x2, [5, 9], axis=1))
    assert np.array_equal(d1[0, ok1], g1[0, ok1]) and w1.shape[1] == 599


def _same_with_nans(got, want):
    gn, wn = np.isnan(got), np.isnan(want)
    return np.array_equal(gn, wn) and got[~gn].tobytes() == want[~wn].tobytes()


@pytest.mark.parametrize("n,m", [(600, 700), (5000, 4500), (8200, 8300)])  # single-launch, two-pass, symmetric
@pytest.mark.parametrize("first_too", [False, True])
def test_nonfinite_inputs_follow_the_reference_rule(cuda_lib, device, n, m, first_too):
    """my_lib.cpp:16 `if (k == 0 || d < best)`: NaN candidates are skipped unless they come first, +inf competes
    normally.  Non-finite points in the middle of both clouds, and (first_too) at index 0 of either cloud."""
    x1 = synthetic.uniform(2, n, 11).numpy()
    x2 = synthetic.uniform(2, m, 12).numpy()
    x1[0, 3] = np.nan
    x1[1, n - 1, 2] = np.inf
    x1[1, 17] = 3e19           # squares overflow to +inf
    x2[0, 5] = np.inf
    x2[0, 9, 1] = np.nan
    x2[1, m // 2] = -np.inf
    if first_too:
        x1[0, 0, 1] = np.nan   # every column of batch 0 sees a NaN first candidate
        x2[1, 0] = np.inf      # every row of batch 1 sees d(k=0) = +inf (or NaN for the inf rows)
        x2[0, 0, 0] = np.nan
    d1, d2, i1, i2 = gpu_forward(x1, x2, device)
    w1, w2, j1, j2 = oracle.forward(x1, x2)
    assert np.array_equal(i1, j1) and np.array_equal(i2, j2)
    assert _same_with_nans(d1, w1) and _same_with_nans(d2, w2)


def test_overflow_to_inf_keeps_index_zero(cuda_lib, device):
    q = np.zeros((1, 3, 3), np.float32)
    t = np.full((1, 40, 3), 3e19, np.float32)  # every squared distance overflows to +inf
    d1, _, i1, _ = gpu_forward(q, t, device)
    assert np.isinf(d1).all() and (i1 == 0).all()  # my_lib.cpp:16: k == 0 is accepted, nothing is < inf


def test_empty_clouds(cuda_lib, device):
    x1 = torch.rand(2, 5, 3, device=device)
    e = torch.zeros(2, 0, 3, device=device)
    d1, d2, i1, i2 = F.nnd_forward_raw(x1, e)
    assert d1.shape == (2, 5) and (d1 == 0).all() and (i1 == 0).all() and d2.shape == (2, 0)
    g1, g2 = F.nnd_backward_raw(x1, e, torch.ones(2, 5, device=device), torch.ones(2, 0, device=device), i1, i2)
    assert (g1 == 0).all() and g2.shape == (2, 0, 3)
    z = torch.zeros(0, 4, 3, device=device)
    out = F.nnd_forward_raw(z, z)
    assert out[0].shape == (0, 4)


# ---------------------------------------------------------------- one-direction entry point
def test_query_entry_point(cuda_lib, device):
    q, t = synthetic.make_pair("crosscopy", 2, 3001, 1777, 42)
    d, i = F.nnd_query_raw(q.to(device), t.to(device))
    wd, wi = oracle.query(q.numpy(), t.numpy(), threads=8)
    assert d.cpu().numpy().tobytes() == wd.tobytes() and i.cpu().numpy().tobytes() == wi.tobytes()


# ---------------------------------------------------------------- backward
@pytest.mark.parametrize("b,n,m", [(1, 1, 1), (2, 100, 37), (1, 5000, 3000), (4, 2048, 2048)])
def test_backward_vs_oracle(b, n, m, cuda_lib, device):
    x1, x2 = synthetic.make_pair("uniform", b, n, m, 43)
    gd1, gd2 = synthetic.grad_pair(b, n, m, 43)
    _, _, i1, i2 = oracle.forward(x1.numpy(), x2.numpy(), threads=8)
    want = oracle.backward(x1.numpy(), x2.numpy(), gd1.numpy(), gd2.numpy(), i1, i2)
    assert_grad_close(gpu_backward(x1, x2, gd1, gd2, i1, i2, device), want)


def test_backward_heavy_collisions(cuda_lib, device):
    """Every query of cloud 1 has the same nearest neighbour: 20000 scatter contributions into one row."""
    n, m = 20000, 64
    x1 = synthetic.uniform(1, n, 44) * 0.01
    x2 = synthetic.uniform(1, m, 45) + 10.0
    x2[0, 17] = 0.005
    gd1, gd2 = synthetic.grad_pair(1, n, m, 44)
    _, _, i1, i2 = oracle.forward(x1.numpy(), x2.numpy(), threads=8)
    assert (i1 == 17).all()
    want = oracle.backward(x1.numpy(), x2.numpy(), gd1.numpy(), gd2.numpy(), i1, i2)
    got = gpu_backward(x1, x2, gd1, gd2, i1, i2, device)
    # 20000 terms summed in a different order: allow 1e-5 of the sum of magnitudes (the conditioning of the sum)
    scale = np.abs(2 * gd1.numpy()[0, :, None] * (x1.numpy()[0] - x2.numpy()[0, 17])).sum(0).max()
    assert np.abs(got[1] - want[1]).max() <= GRAD_RTOL * scale
    assert_grad_close((got[0],), (want[0],))


def test_backward_overwrites_garbage(cuda_lib, device):
    """The C ABI zero-initialises the gradients itself (my_lib.cpp:91-94), whatever the buffers held."""
    import ctypes

    x1, x2 = synthetic.make_pair("uniform", 1, 300, 200, 46)
    gd1, gd2 = synthetic.grad_pair(1, 300, 200, 46)
    _, _, i1, i2 = oracle.forward(x1.numpy(), x2.numpy())
    want = oracle.backward(x1.numpy(), x2.numpy(), gd1.numpy(), gd2.numpy(), i1, i2)
    t = [a.to(device) for a in (x1, x2, gd1, gd2, torch.from_numpy(i1), torch.from_numpy(i2))]
    g1 = torch.full((1, 300, 3), 7.0, device=device)
    g2 = torch.full((1, 200, 3), float("nan"), device=device)
    rc = cuda_lib.nnd_backward_cuda_f32(t[0].data_ptr(), t[1].data_ptr(), t[2].data_ptr(), t[3].data_ptr(),
                                        t[4].data_ptr(), t[5].data_ptr(), 1, 300, 200, g1.data_ptr(), g2.data_ptr(),
                                        ctypes.c_void_p(torch.cuda.current_stream().cuda_stream))
    assert rc == 0
    assert_grad_close((g1.cpu().numpy(), g2.cpu().numpy()), want)


# ---------------------------------------------------------------- autograd boundary (reference test.py semantics)
def test_reference_smoke_script_semantics(cuda_lib, device):
    """dip/torch-nndistance/test.py:8-29 with asserts: non-leaf then leaf inputs, loss = sum(dist1)."""
    import torch_nndistance as NND

    p1, p2 = synthetic.make_pair("uniform", 16, 2048, 1024, 21)
    want_fwd = oracle.forward(p1.numpy(), p2.numpy(), threads=8)
    ones, zeros = np.ones((16, 2048), np.float32), np.zeros((16, 1024), np.float32)
    want_g1, want_g2 = oracle.backward(p1.numpy(), p2.numpy(), ones, zeros, want_fwd[2], want_fwd[3])

    points1 = p1.clone().requires_grad_(True)  # leaf on the CPU, .cuda() makes a non-leaf (test.py:10-13)
    c1, c2 = points1.to(device), p2.to(device)
    dist1, dist2 = NND.nnd(c1, c2)
    assert dist1.detach().cpu().numpy().tobytes() == want_fwd[0].tobytes()
    assert dist2.detach().cpu().numpy().tobytes() == want_fwd[1].tobytes()
    torch.sum(dist1).backward()
    assert_grad_close((points1.grad.numpy(),), (want_g1,))

    l1 = p1.to(device).requires_grad_(True)  # leaf on the GPU (test.py:22-23)
    l2 = p2.to(device).requires_grad_(True)
    d1, d2 = NND.nnd(l1, l2)
    torch.sum(d1).backward()
    assert_grad_close((l1.grad.cpu().numpy(), l2.grad.cpu().numpy()), (want_g1, want_g2))


def test_nnd_with_idx_and_module(cuda_lib, device):
    import torch_nndistance as NND

    x1, x2 = synthetic.make_pair("lattice", 2, 900, 1100, 47)
    a = x1.to(device).requires_grad_(True)
    d1, d2, i1, i2 = NND.nnd_with_idx(a, x2.to(device))
    assert i1.dtype == torch.int32 and not i1.requires_grad and d1.requires_grad
    want = oracle.forward(x1.numpy(), x2.numpy(), threads=8)
    assert_forward_exact(tuple(t.detach().cpu().numpy() for t in (d1, d2, i1, i2)), want)
    (d1.mean() + d2.mean()).backward()  # chamfer loss as the callers form it (dip/train.py:84)
    gd = (np.full((2, 900), 1 / 1800, np.float32), np.full((2, 1100), 1 / 2200, np.float32))
    wg = oracle.backward(x1.numpy(), x2.numpy(), gd[0], gd[1], want[2], want[3])
    assert_grad_close((a.grad.cpu().numpy(),), (wg[0],))
    m1, m2 = NND.NNDModule()(x1.to(device), x2.to(device))
    assert torch.equal(m1, d1.detach()) and torch.equal(m2, d2.detach())


def test_aten_shim_writes_in_place(cuda_lib, device):
    import torch_nndistance_aten as my_lib

    x1, x2 = synthetic.make_pair("uniform", 2, 333, 444, 48)
    a, b = x1.to(device), x2.to(device)
    dist1 = torch.zeros(2, 333, device=device)
    dist2 = torch.zeros(2, 444, device=device)
    idx1 = torch.zeros(2, 333, dtype=torch.int32, device=device)
    idx2 = torch.zeros(2, 444, dtype=torch.int32, device=device)
    assert my_lib.nnd_forward_cuda(a, b, dist1, dist2, idx1, idx2) == 1  # nnd_cuda.cu:161
    want = oracle.forward(x1.numpy(), x2.numpy())
    assert_forward_exact(tuple(t.cpu().numpy() for t in (dist1, dist2, idx1, idx2)), want)
    gd1, gd2 = synthetic.grad_pair(2, 333, 444, 48)
    g1, g2 = torch.zeros_like(a), torch.zeros_like(b)
    assert my_lib.nnd_backward_cuda(a, b, g1, g2, gd1.to(device), gd2.to(device), idx1, idx2) == 1
    assert_grad_close((g1.cpu().numpy(), g2.cpu().numpy()),
                      oracle.backward(x1.numpy(), x2.numpy(), gd1.numpy(), gd2.numpy(), want[2], want[3]))
    with pytest.raises(RuntimeError, match="contiguous"):
        my_lib.nnd_forward_cuda(a[:, ::2], b, dist1, dist2, idx1, idx2)  # my_lib_cuda.cpp:5


# ---------------------------------------------------------------- the panel scan (columns in uniform registers)
@pytest.fixture
def panels_forced(cuda_lib):
    """Push every symmetric scan through nnd_scan_symc_kernel (normally only row ranges >= 200 000 rows)."""
    before = cuda_lib.nnd_set_panel_min_rows(1)
    small = cuda_lib.nnd_set_small_max_points(0)  # a lone pair of <= 10 k points would take the single-launch kernel
    yield
    cuda_lib.nnd_set_panel_min_rows(before)
    cuda_lib.nnd_set_small_max_points(small)


@pytest.mark.parametrize("kind,b,n,m", [
    ("uniform", 1, 8192, 8192),        # exactly 4 panels of 4 tiles
    ("uniform", 2, 9001, 10241),       # ragged rows, ragged last panel (21 tiles), two batch elements
    ("clustered", 1, 20000, 8200),     # last panel = 1 tile + 8 points
    ("lattice", 1, 12000, 12345),      # exact ties everywhere: lowest index must win through the panels too
    ("identical", 1, 9000, 9000),
    ("crosscopy", 3, 8192, 16000),
])
def test_panel_scan_matches_oracle(panels_forced, cuda_lib, device, kind, b, n, m):
    x1, x2 = synthetic.make_pair(kind, b, n, m, 43)
    want = oracle.forward(x1.numpy(), x2.numpy(), threads=8)
    assert cuda_lib.nnd_forward_is_symmetric(n, m)
    got = gpu_forward(x1, x2, device)
    assert_forward_exact(got, want, what=f"panel {kind}: ")
    assert cuda_lib.nnd_last_launch_count() >= 3 + (m + 2047) // 2048 * b  # one launch per panel (at most 4 tiles) at least


@pytest.mark.parametrize("tiles", [1, 2, 4])
def test_panel_scan_equals_persistent_scan_bitwise(cuda_lib, device, tiles):
    """Same inputs through both scan kernels: all four outputs identical bit for bit (262144 points, clustered)."""
    x1, x2 = synthetic.make_pair("clustered", 1, 262144, 200000, 44)
    before = cuda_lib.nnd_set_panel_min_rows(1 << 40)
    try:
        ref = gpu_forward(x1, x2, device)
        cuda_lib.nnd_set_panel_min_rows(1)
        cuda_lib.nnd_set_panel_tiles(tiles)
        got = gpu_forward(x1, x2, device)
        assert cuda_lib.nnd_last_launch_count() >= 200000 // (512 * tiles)
    finally:
        cuda_lib.nnd_set_panel_min_rows(before)
        cuda_lib.nnd_set_panel_tiles(0)
    for a, c in zip(got, ref):
        assert a.tobytes() == c.tobytes()


def test_panel_scan_under_graph_capture_and_side_stream(panels_forced, cuda_lib, device):
    """The panel path forks into internal streams and joins back: it must behave like one operation on the caller's
    stream.  While that stream is being captured into a CUDA graph the forward takes the persistent kernel instead
    (the panel scan's per-device state cannot be ordered against graph replays); same bits either way."""
    x1, x2 = synthetic.make_pair("uniform", 1, 10000, 9000, 45)
    want = oracle.forward(x1.numpy(), x2.numpy(), threads=8)
    a, c = x1.to(device), x2.to(device)
    s = torch.cuda.Stream(device=device)
    with torch.cuda.stream(s):
        out = F.nnd_forward_raw(a, c)
    s.synchronize()
    assert_forward_exact(tuple(t.cpu().numpy() for t in out), want, what="side stream: ")
    g = torch.cuda.CUDAGraph()
    static_a = a.clone()
    torch.cuda.synchronize()
    with torch.cuda.graph(g):
        cap = F.nnd_forward_raw(static_a, c)
    static_a.copy_(a.flip(1))  # new rows, replay must see them
    g.replay()
    torch.cuda.synchronize()
    want2 = oracle.forward(x1.flip(1).contiguous().numpy(), x2.numpy(), threads=8)
    assert_forward_exact(tuple(t.cpu().numpy() for t in cap), want2, what="graph replay: ")


def test_captured_large_forward_replays_while_an_eager_large_forward_runs(panels_forced, cuda_lib, device):
    """a captured graph that contained a panel scan could corrupt results when replayed while
    another large forward ran.  Now a capture never contains a panel scan.  Replay a captured large forward on one
    stream while eager large forwards (panel scans) run on another, never synchronising the two: both bit-exact."""
    xa1, xa2 = synthetic.make_pair("uniform", 1, 30000, 24000, 56)
    xb1, xb2 = synthetic.make_pair("clustered", 1, 20000, 26000, 57)
    want_a = oracle.forward(xa1.numpy(), xa2.numpy(), threads=16)
    want_b = oracle.forward(xb1.numpy(), xb2.numpy(), threads=16)
    a1, a2, b1, b2 = (t.to(device) for t in (xa1, xa2, xb1, xb2))
    sg, se = torch.cuda.Stream(device=device), torch.cuda.Stream(device=device)
    g = torch.cuda.CUDAGraph()
    torch.cuda.synchronize()
    with torch.cuda.graph(g, stream=sg):
        cap = F.nnd_forward_raw(a1, a2)
    outs_b, snaps = [], []
    for _ in range(5):
        with torch.cuda.stream(sg):
            g.replay()
            snaps.append(tuple(t.clone() for t in cap))
        with torch.cuda.stream(se):
            outs_b.append(F.nnd_forward_raw(b1, b2))
            assert cuda_lib.nnd_last_launch_count() > 10  # eager: the panel scan (one launch per panel)
    torch.cuda.synchronize()
    for snap in snaps:
        assert_forward_exact(tuple(t.cpu().numpy() for t in snap), want_a, what="graph replay: ")
    for out in outs_b:
        assert_forward_exact(tuple(t.cpu().numpy() for t in out), want_b, what="eager beside the replay: ")


def test_panel_scans_from_two_streams_with_different_panel_sizes(panels_forced, cuda_lib, device):
    """Panel scans share one constant window per device.  Two caller streams that alternate scans with different
    panel sizes (the window is cut 2 x 4 tiles for one, 8 x 1 tile for the other) and never synchronise with each
    other must still get the single-stream results."""
    xa1, xa2 = synthetic.make_pair("uniform", 1, 30000, 24000, 46)
    xb1, xb2 = synthetic.make_pair("clustered", 1, 16384, 20000, 47)
    a1, a2, b1, b2 = (t.to(device) for t in (xa1, xa2, xb1, xb2))
    try:
        cuda_lib.nnd_set_panel_tiles(4)
        want_a = tuple(t.cpu() for t in F.nnd_forward_raw(a1, a2))
        cuda_lib.nnd_set_panel_tiles(1)
        want_b = tuple(t.cpu() for t in F.nnd_forward_raw(b1, b2))
        torch.cuda.synchronize()
        sa, sb = torch.cuda.Stream(device=device), torch.cuda.Stream(device=device)
        outs_a, outs_b = [], []
        for _ in range(6):
            cuda_lib.nnd_set_panel_tiles(4)
            with torch.cuda.stream(sa):
                outs_a.append(F.nnd_forward_raw(a1, a2))
            cuda_lib.nnd_set_panel_tiles(1)
            with torch.cuda.stream(sb):
                outs_b.append(F.nnd_forward_raw(b1, b2))
        torch.cuda.synchronize()
    finally:
        cuda_lib.nnd_set_panel_tiles(0)
    for outs, want in ((outs_a, want_a), (outs_b, want_b)):
        for out in outs:
            for got, w in zip(out, want):
                assert torch.equal(got.cpu(), w)


def test_non_contiguous_inputs_and_side_stream(cuda_lib, device):
    import torch_nndistance as NND

    x1 = torch.rand(2, 3, 700, generator=torch.Generator().manual_seed(1)).transpose(1, 2)  # [2,700,3] strided, as DIP makes it (dip/train.py:84)
    x2 = torch.rand(2, 3, 650, generator=torch.Generator().manual_seed(2)).transpose(1, 2)
    want = oracle.forward(x1.contiguous().numpy(), x2.contiguous().numpy())
    s = torch.cuda.Stream(device=device)
    a, b = x1.to(device), x2.to(device)
    assert not a.is_contiguous()
    torch.cuda.synchronize()
    with torch.cuda.stream(s):
        d1, d2 = NND.nnd(a, b)
    s.synchronize()
    assert d1.detach().cpu().numpy().tobytes() == want[0].tobytes()
    assert d2.detach().cpu().numpy().tobytes() == want[1].tobytes()


# ---------------------------------------------------------------- full-size configs: sampled-row exactness + properties
def _sampled_rows_exact(x1, x2, got, rows1, rows2):
    """Exact check of a sample of rows of each direction against the oracle (full target clouds)."""
    d1, d2, i1, i2 = got
    for b in range(x1.shape[0]):
        wd, wi = oracle.query(x1[b:b + 1, rows1], x2[b:b + 1], threads=16)
        assert d1[b, rows1].tobytes() == wd[0].tobytes() and i1[b, rows1].tobytes() == wi[0].tobytes()
        wd, wi = oracle.query(x2[b:b + 1, rows2], x1[b:b + 1], threads=16)
        assert d2[b, rows2].tobytes() == wd[0].tobytes() and i2[b, rows2].tobytes() == wi[0].tobytes()


def _forward_properties(x1, x2, got):
    """Size-independent properties: the reported neighbour really is at the reported distance (same
    rounding), and no reported distance beats a random probe of other candidates."""
    d1, d2, i1, i2 = got
    b = 0
    for q, t, d, i in ((x1[b], x2[b], d1[b], i1[b]), (x2[b], x1[b], d2[b], i2[b])):
        assert ((i >= 0) & (i < t.shape[0])).all()
        diff = t[i] - q
        sq = diff * diff
        assert ((sq[:, 0] + sq[:, 1]) + sq[:, 2]).astype(np.float32).tobytes() == d.tobytes()
        probe = np.random.default_rng(0).integers(0, t.shape[0], size=q.shape[0])
        dp = t[probe] - q
        sp = dp * dp
        alt = (sp[:, 0] + sp[:, 1]) + sp[:, 2]
        assert (d <= alt).all()
        assert not ((alt == d) & (probe < i)).any()  # an equal candidate with a lower index would be a tie-break bug


def _assert_full_digest(name, got, x1, x2, device, grads=True):
    """WHOLE output tensors against the SHA-256 digests the compiled reference produced for this config
    (tests/golden/digests_full.json, made by tests/golden/make_golden_full.py): every row, both directions, and the
    deterministic backward's gradients bit for bit."""
    meta = FULL_DIGESTS[name]
    assert sha(x1.numpy()) == meta["input_sha256"]["xyz1"] and sha(x2.numpy()) == meta["input_sha256"]["xyz2"], \
        f"{name}: regenerated inputs differ from the golden run"
    for k, v in zip(("dist1", "dist2", "idx1", "idx2"), got):
        assert sha(v) == meta["sha256"][k], f"{name}: {k} differs from the reference's full-size output"
    if grads:
        b, n, m = meta["b"], meta["n"], meta["m"]
        gd1, gd2 = synthetic.grad_pair(b, n, m, meta["config"])
        g1, g2 = gpu_backward(x1, x2, gd1, gd2, got[2], got[3], device, exact=True)
        assert sha(g1) == meta["sha256"]["grad1"] and sha(g2) == meta["sha256"]["grad2"], \
            f"{name}: deterministic backward differs from the reference's full-size gradients"


@pytest.mark.parametrize("kind", ["lattice", "duplicates", "identical", "crosscopy", "clustered"])
def test_config5_tie_stress_262144(kind, cuda_lib, device):
    """BASELINE.json configs[4]: N = M = 262144 degenerate clouds: all 4 x 262144 outputs vs the reference digest."""
    n = 262144
    x1, x2 = synthetic.make_pair(kind, 1, n, n, 5)
    got = gpu_forward(x1, x2, device)
    assert cuda_lib.nnd_last_launch_count() > 100  # the panel scan, the kernel that produces the headline number
    _assert_full_digest(f"c5_{kind}", got, x1, x2, device)
    x1n, x2n = x1.numpy(), x2.numpy()
    _forward_properties(x1n, x2n, got)
    rows = np.random.default_rng(1).choice(n, 1024, replace=False)
    rows.sort()
    _sampled_rows_exact(x1n, x2n, got, rows, rows)
    if kind == "identical":
        assert (got[0] == 0).all() and (got[1] == 0).all()
        h = n // 2
        want = np.concatenate([np.arange(h), np.arange(h)]).astype(np.int32)  # lowest duplicate index
        assert np.array_equal(got[2][0], want) and np.array_equal(got[3][0], want)


def test_config2_dip_training_shape(cuda_lib, device):
    """BASELINE.json configs[1]: B=32, N=M=16384 (sampled rows exact + properties on batch 0)."""
    x1, x2 = synthetic.make_pair("uniform", 32, 16384, 16384, 2)
    got = gpu_forward(x1, x2, device)
    _assert_full_digest("c2_uniform", got, x1, x2, device)
    _forward_properties(x1.numpy(), x2.numpy(), got)
    rows = np.arange(0, 16384, 257)
    _sampled_rows_exact(x1.numpy()[[0, 17, 31]], x2.numpy()[[0, 17, 31]], tuple(g[[0, 17, 31]] for g in got), rows, rows)


@pytest.mark.parametrize("kind", ["uniform", "clustered"])
def test_config4_one_million(kind, cuda_lib, device):
    """BASELINE.json configs[3] on one GPU: N = M = 1,000,000."""
    n = 1_000_000
    x1, x2 = synthetic.make_pair(kind, 1, n, n, 4)
    got = gpu_forward(x1, x2, device)
    _assert_full_digest(f"c4_{kind}", got, x1, x2, device, grads=False)  # gradients: below, with the bench's inputs
    _forward_properties(x1.numpy(), x2.numpy(), got)
    rows = np.random.default_rng(2).choice(n, 512, replace=False)
    rows.sort()
    _sampled_rows_exact(x1.numpy(), x2.numpy(), got, rows, rows)
    # backward at full size: linearity (grad(2g) == 2 grad(g) exactly is not guaranteed with atomics; use the
    # conservation property instead: every contribution is added to one cloud and subtracted from the other)
    gd1, gd2 = synthetic.grad_pair(1, n, n, 4, mean=True)
    g1, g2 = gpu_backward(x1, x2, gd1, gd2, got[2], got[3], device)
    tot = g1.astype(np.float64).sum(axis=(0, 1)) + g2.astype(np.float64).sum(axis=(0, 1))
    mag = np.abs(g1).astype(np.float64).sum() + np.abs(g2).astype(np.float64).sum()
    assert np.abs(tot).max() <= 1e-5 * mag
    wg = oracle.backward(x1.numpy(), x2.numpy(), gd1.numpy(), gd2.numpy(), got[2], got[3])
    assert_grad_close((g1, g2), wg)
    ge = gpu_backward(x1, x2, gd1, gd2, got[2], got[3], device, exact=True)
    assert_grad_bits(ge, wg, what=f"1M {kind} (exact): ")
    meta = FULL_DIGESTS[f"c4_{kind}"]["sha256"]  # the compiled reference's own gradients for the same inputs
    assert sha(ge[0]) == meta["grad1_mean"] and sha(ge[1]) == meta["grad2_mean"]


def test_config3_ndp_loop_100k(cuda_lib, device):
    """BASELINE.json configs[2]: 500 iterations of chamfer on a 100k pair where cloud 1 moves down its own gradient
    (x <- x - 0.01 grad, with grad of the SUM so the step is visible).  Iteration 0 is compared with the
    reference's full-size digest; every 50th iteration (and the last) is compared with the oracle on sampled rows."""
    import torch_nndistance as NND

    n = 100_000
    x1, x2 = synthetic.make_pair("uniform", 1, n, n, 3)
    a = x1.to(device).requires_grad_(True)
    bt = x2.to(device)
    rows = np.arange(0, n, 199)
    losses = []
    for it in range(500):
        d1, d2, i1, i2 = NND.nnd_with_idx(a, bt)
        loss = d1.mean() + d2.mean()
        (g,) = torch.autograd.grad(loss, a)
        if it == 0:
            _assert_full_digest("c3_uniform", tuple(t.detach().cpu().numpy() for t in (d1, d2, i1, i2)), x1, x2, device)
        if it % 50 == 0 or it == 499:
            an = a.detach().cpu().numpy()
            _sampled_rows_exact(an, x2.numpy(), tuple(t.detach().cpu().numpy() for t in (d1, d2, i1, i2)), rows, rows)
            losses.append(loss.item())
        a = (a.detach() - 0.01 * n * g).requires_grad_(True)
    assert losses[1] < losses[0] and losses[-1] < losses[1]


# ---------------------------------------------------------------- warm start (nnd_forward_warm_cuda_f32)
@pytest.mark.parametrize("kind,b,n,m,panels", [("uniform", 2, 9000, 8200, False), ("lattice", 1, 20000, 30000, False),
                                               ("clustered", 2, 9001, 10241, True), ("identical", 1, 9000, 9000, True)])
def test_warm_start_never_changes_the_results(kind, b, n, m, panels, cuda_lib, device):
    """hint_idx2 only seeds the column filter: whatever it holds -- the true neighbours, neighbours of a perturbed cloud,
    zeros, garbage, out-of-range values -- all four outputs keep the oracle's bits."""
    before = cuda_lib.nnd_set_panel_min_rows(1) if panels else None
    small = cuda_lib.nnd_set_small_max_points(0)
    try:
        x1, x2 = synthetic.make_pair(kind, b, n, m, 58)
        want = oracle.forward(x1.numpy(), x2.numpy(), threads=16)
        a, c = x1.to(device), x2.to(device)
        cold = F.nnd_forward_raw(a, c)
        assert_forward_exact(tuple(t.cpu().numpy() for t in cold), want, what="cold: ")
        moved = F.nnd_forward_raw(a + 0.01 * torch.randn_like(a), c)[3]
        g = torch.Generator().manual_seed(3)
        hints = {"true": cold[3], "previous iteration": moved, "zeros": torch.zeros_like(cold[3]),
                 "random": torch.randint(0, n, (b, m), generator=g, dtype=torch.int32).to(device),
                 "out of range": torch.randint(-5 * n, 5 * n, (b, m), generator=g, dtype=torch.int32).to(device)}
        for name, h in hints.items():
            got = F.nnd_forward_raw(a, c, hint_idx2=h)
            assert_forward_exact(tuple(t.cpu().numpy() for t in got), want, what=f"hint = {name}: ")
        aliased = cold[3].clone()  # the loop's own buffer as hint AND output is what ndp_loop does through the C ABI
        with pytest.raises(ValueError):
            F.nnd_forward_raw(a, c, hint_idx2=aliased.long())
    finally:
        cuda_lib.nnd_set_small_max_points(small)
        if panels:
            cuda_lib.nnd_set_panel_min_rows(before)


@pytest.mark.parametrize("kind,b,n,m", [("uniform", 1, 9000, 8200), ("clustered", 2, 9001, 10241), ("lattice", 1, 12000, 12345),
                                        ("identical", 1, 9000, 9000), ("large", 1, 8200, 9000), ("tiny", 1, 8200, 9000)])
def test_automatic_hints_never_change_the_results(kind, b, n, m, cuda_lib, device):
    """The hash-grid hints (csrc/nnd_hint.cuh) only seed the column filter: off, in front of the panel scan, in front of
    the persistent kernel -- and with non-finite points in the clouds -- the outputs keep the oracle's bits."""
    small = cuda_lib.nnd_set_small_max_points(0)
    keep = cuda_lib.nnd_set_auto_hint(-1)
    try:
        x1, x2 = synthetic.make_pair(kind, b, n, m, 59)
        x1, x2 = x1.clone(), x2.clone()
        if kind == "uniform":
            x1[0, 7] = float("inf")
            x2[0, 11, 1] = float("nan")
            x1[0, 100] = 3e30
        want = oracle.forward(x1.numpy(), x2.numpy(), threads=16)
        for panels in (False, True):
            before = cuda_lib.nnd_set_panel_min_rows(1 if panels else 1 << 40)
            try:
                for mode in (0, 1, 2):
                    cuda_lib.nnd_set_auto_hint(mode)
                    d1, d2, i1, i2 = gpu_forward(x1, x2, device)
                    assert np.array_equal(i1, want[2]) and np.array_equal(i2, want[3]), (panels, mode)
                    assert _same_with_nans(d1, want[0]) and _same_with_nans(d2, want[1]), (panels, mode)
            finally:
                cuda_lib.nnd_set_panel_min_rows(before)
    finally:
        cuda_lib.nnd_set_auto_hint(keep)
        cuda_lib.nnd_set_small_max_points(small)
