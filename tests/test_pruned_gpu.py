"""GPU parity tests of the exact pruned search (nnd_forward_pruned_cuda_f32, csrc/nnd_pruned.cuh): it must return the
SAME BITS as the brute-force scan of the reference (my_lib.cpp:3-25) -- dist bit-exact, lowest index on ties, the
NaN / inf rules of my_lib.cpp:16 -- on every kind of cloud.  Checked against the CPU oracle at sizes it finishes in
seconds and against the compiled reference's full-size digests (tests/golden/digests_full.json) at C5 / C4."""
import numpy as np
import pytest
import torch

import oracle
from _util import FULL_DIGESTS, assert_forward_exact, sha
from pointcloudregistration_b200 import functional as F
from pointcloudregistration_b200 import synthetic

pytestmark = pytest.mark.gpu


def pruned(x1, x2, device):
    out = F.nnd_forward_pruned_raw(torch.as_tensor(x1).to(device), torch.as_tensor(x2).to(device))
    torch.cuda.synchronize()
    return tuple(t.cpu().numpy() for t in out)


def _same_with_nans(got, want):
    gn, wn = np.isnan(got), np.isnan(want)
    return np.array_equal(gn, wn) and got[~gn].tobytes() == want[~wn].tobytes()


@pytest.mark.parametrize("kind", synthetic.KINDS)
@pytest.mark.parametrize("b,n,m", [(2, 3000, 2500), (1, 20000, 17001), (3, 65, 700)])
def test_pruned_matches_the_oracle(kind, b, n, m, cuda_lib, device):
    if kind == "identical":
        m = n
    x1, x2 = synthetic.make_pair(kind, b, n, m, 7)
    got = pruned(x1, x2, device)
    want = oracle.forward(x1.numpy(), x2.numpy(), threads=8)
    assert_forward_exact(got, want, what=f"pruned {kind} {b}x{n}x{m}: ")


@pytest.mark.parametrize("b,n,m", [(1, 1, 1), (2, 1, 300), (2, 300, 1), (1, 5, 70), (4, 64, 64), (1, 63, 65), (1, 2049, 33)])
def test_pruned_tiny_and_ragged_tile_shapes(b, n, m, cuda_lib, device):
    x1, x2 = synthetic.make_pair("lattice", b, n, m, 8)  # ties everywhere
    got = pruned(x1, x2, device)
    assert_forward_exact(got, oracle.forward(x1.numpy(), x2.numpy()), what=f"pruned {b}x{n}x{m}: ")


def test_pruned_random_shapes_on_coarse_grids(cuda_lib, device):
    """Random shapes, coordinates on coarse grids of random pitch (exact ties and duplicates everywhere, the lowest
    ORIGINAL index must win although the search visits targets in curve order), random offsets and scales."""
    rng = np.random.default_rng(20261105)
    for case in range(24):
        b = int(rng.integers(1, 4))
        n, m = int(rng.integers(1, 3000)), int(rng.integers(1, 3000))
        cells = int(rng.choice([2, 3, 7, 16, 64, 1000]))
        scale = float(rng.choice([1.0, 1e-3, 37.5, 4096.0]))
        off = rng.normal(size=3).astype(np.float32) * np.float32(scale)
        x1 = (rng.integers(0, cells, size=(b, n, 3)).astype(np.float32) / np.float32(cells)) * np.float32(scale) + off
        x2 = (rng.integers(0, cells, size=(b, m, 3)).astype(np.float32) / np.float32(cells)) * np.float32(scale) + off
        got = pruned(x1, x2, device)
        assert_forward_exact(got, oracle.forward(x1, x2), what=f"case {case} ({b}x{n}x{m}, {cells} cells, scale {scale}): ")


def test_pruned_query_one_direction(cuda_lib, device):
    """nnd_query_pruned_cuda_f32 through functional.nnd_query_raw with the search mode switched on."""
    q, t = synthetic.make_pair("clustered", 1, 70000, 50000, 10)  # 3.5e9 evaluations: above the routing threshold
    qd, td = q.to(device), t.to(device)
    want = F.nnd_query_raw(qd, td)
    brute_launches = cuda_lib.nnd_last_launch_count()
    before = F.set_search_mode("pruned")
    try:
        got = F.nnd_query_raw(qd, td)
        assert cuda_lib.nnd_last_launch_count() == 11 != brute_launches  # 2 boxes + 2 x (code, gather, 2 box levels) + 1 search
    finally:
        F.set_search_mode(before)
    assert torch.equal(got[0], want[0]) and torch.equal(got[1], want[1])
    w = oracle.query(q.numpy()[:, :3000], t.numpy(), threads=8)
    assert got[0].cpu().numpy()[:, :3000].tobytes() == w[0].tobytes() and got[1].cpu().numpy()[:, :3000].tobytes() == w[1].tobytes()


@pytest.mark.parametrize("share", [1, 2])  # one query group per warp / one per CTA, whatever the size
def test_pruned_both_group_mappings(share, cuda_lib, device):
    """Every shape class through BOTH search kernels (nnd_set_pruned_share), incl. ties and non-finite points."""
    before = cuda_lib.nnd_set_pruned_share(share)
    try:
        for kind, b, n, m in (("lattice", 2, 3000, 2500), ("crosscopy", 1, 20000, 17001), ("duplicates", 3, 65, 700),
                              ("clustered", 1, 33, 40000), ("tiny", 2, 700, 900)):
            x1, x2 = synthetic.make_pair(kind, b, n, m, 12)
            got = pruned(x1, x2, device)
            assert_forward_exact(got, oracle.forward(x1.numpy(), x2.numpy(), threads=8), what=f"share {share} {kind}: ")
        rng = np.random.default_rng(7 + share)
        for case in range(6):  # NaN / +-inf / overflowing values sprinkled at random, sometimes at index 0
            b, n, m = int(rng.integers(1, 3)), int(rng.integers(40, 4000)), int(rng.integers(40, 4000))
            x1 = rng.random((b, n, 3), dtype=np.float32)
            x2 = rng.random((b, m, 3), dtype=np.float32)
            for x in (x1, x2):
                k = int(rng.integers(1, 12))
                pos = rng.integers(0, x.shape[1], size=k)
                pos[0] = 0 if case % 2 else pos[0]
                x[rng.integers(0, b, size=k), pos, rng.integers(0, 3, size=k)] = rng.choice(
                    np.array([np.nan, np.inf, -np.inf, 3e19, -3e19], np.float32), size=k)
            d1, d2, i1, i2 = pruned(x1, x2, device)
            w1, w2, j1, j2 = oracle.forward(x1, x2)
            assert np.array_equal(i1, j1) and np.array_equal(i2, j2), (share, case)
            assert _same_with_nans(d1, w1) and _same_with_nans(d2, w2), (share, case)
    finally:
        cuda_lib.nnd_set_pruned_share(before)


def test_pruned_workspace_errors(cuda_lib, device):
    x1, x2 = synthetic.make_pair("uniform", 1, 5000, 4000, 3)
    a, c = x1.to(device), x2.to(device)
    outs = [torch.empty(1, 5000, device=device), torch.empty(1, 4000, device=device),
            torch.empty(1, 5000, dtype=torch.int32, device=device), torch.empty(1, 4000, dtype=torch.int32, device=device)]
    need = cuda_lib.nnd_pruned_workspace_bytes(1, 5000, 4000)
    ws = torch.empty(need + 256, dtype=torch.uint8, device=device)
    args = (a.data_ptr(), c.data_ptr(), 1, 5000, 4000, outs[0].data_ptr(), outs[1].data_ptr(), outs[2].data_ptr(), outs[3].data_ptr())
    assert cuda_lib.nnd_forward_pruned_cuda_f32(*args, ws.data_ptr(), need - 1, None) == 10002       # too small
    assert cuda_lib.nnd_forward_pruned_cuda_f32(*args, ws.data_ptr() + 16, need, None) == 10002      # misaligned
    assert cuda_lib.nnd_forward_pruned_reuse_cuda_f32(*args, ws.data_ptr(), need, 4, None) == 10001  # unknown reuse bits
    assert cuda_lib.nnd_forward_pruned_cuda_f32(*args, ws.data_ptr(), need, None) == 0
    torch.cuda.synchronize()
    want = oracle.forward(x1.numpy(), x2.numpy())
    assert_forward_exact(tuple(t.cpu().numpy() for t in outs), want)


def test_pruned_empty_clouds(cuda_lib, device):
    x1 = synthetic.uniform(2, 40, 1)
    e = torch.zeros(2, 0, 3)
    d1, d2, i1, i2 = pruned(x1, e, device)  # my_lib.cpp:9-10,21-22: an empty inner loop leaves dist = 0, idx = 0
    assert d1.shape == (2, 40) and d2.shape == (2, 0) and not d1.any() and not i1.any()
    d1, d2, i1, i2 = pruned(e, x1, device)
    assert d2.shape == (2, 40) and not d2.any() and not i2.any()


@pytest.mark.parametrize("n,m", [(600, 700), (9000, 8300)])
@pytest.mark.parametrize("first_too", [False, True])
def test_pruned_nonfinite_inputs_follow_the_reference_rule(cuda_lib, device, n, m, first_too):
    """my_lib.cpp:16 `if (k == 0 || d < best)`: NaN candidates are skipped unless they come first, +inf competes
    normally -- the same inputs as test_nnd_gpu.py's brute-force test."""
    x1 = synthetic.uniform(2, n, 11).numpy()
    x2 = synthetic.uniform(2, m, 12).numpy()
    x1[0, 3] = np.nan
    x1[1, n - 1, 2] = np.inf
    x1[1, 17] = 3e19           # squares overflow to +inf
    x2[0, 5] = np.inf
    x2[0, 9, 1] = np.nan
    x2[1, m // 2] = -np.inf
    if first_too:
        x1[0, 0, 1] = np.nan
        x2[1, 0] = np.inf
        x2[0, 0, 0] = np.nan
    d1, d2, i1, i2 = pruned(x1, x2, device)
    w1, w2, j1, j2 = oracle.forward(x1, x2)
    assert np.array_equal(i1, j1) and np.array_equal(i2, j2)
    assert _same_with_nans(d1, w1) and _same_with_nans(d2, w2)


def test_pruned_overflow_to_inf_keeps_index_zero(cuda_lib, device):
    q = np.zeros((1, 3, 3), np.float32)
    t = np.full((1, 40, 3), 3e19, np.float32)  # every squared distance overflows to +inf
    d1, d2, i1, i2 = pruned(q, t, device)
    w = oracle.forward(q, t)
    assert np.isinf(d1).all() and (i1 == 0).all()
    assert_forward_exact((d1, d2, i1, i2), w)


def test_pruned_all_points_identical(cuda_lib, device):
    """Nothing can be pruned (every bound is 0 = every best): still the lowest index everywhere."""
    x1 = np.full((1, 3000, 3), 0.25, np.float32)
    x2 = np.full((1, 2000, 3), 0.25, np.float32)
    d1, d2, i1, i2 = pruned(x1, x2, device)
    assert not d1.any() and not d2.any() and not i1.any() and not i2.any()


def test_search_mode_switch_routes_the_public_api(cuda_lib, device):
    import torch_nndistance as NND

    x1, x2 = synthetic.make_pair("crosscopy", 2, 40000, 30000, 9)  # 2.4e9 evaluations: above the routing threshold
    a, c = x1.to(device).requires_grad_(True), x2.to(device).requires_grad_(True)
    want = NND.nnd_with_idx(a, c)
    before = F.set_search_mode("pruned")
    try:
        assert F.search_mode() == "pruned"
        got = NND.nnd_with_idx(a, c)
        assert cuda_lib.nnd_last_launch_count() == 12  # 2 boxes + 2 x (code, gather, 2 box levels) + 2 searches
        (got[0].mean() + got[1].mean()).backward()
        small = NND.nnd_with_idx(a[:, :500], c[:, :400])  # below the thresholds: the single-launch kernels
        assert cuda_lib.nnd_last_launch_count() == 1
    finally:
        F.set_search_mode(before)
    assert F.search_mode() == "brute"
    for g, w in zip(got, want):
        assert torch.equal(g, w)
    assert torch.isfinite(a.grad).all() and small[0].shape == (2, 500)
    with pytest.raises(ValueError):
        F.set_search_mode("fast")


@pytest.mark.parametrize("name,kind,n,cfg", [("c5_lattice", "lattice", 262144, 5), ("c5_duplicates", "duplicates", 262144, 5),
                                            ("c5_identical", "identical", 262144, 5), ("c5_crosscopy", "crosscopy", 262144, 5),
                                            ("c5_clustered", "clustered", 262144, 5), ("c3_uniform", "uniform", 100_000, 3),
                                            ("c4_uniform", "uniform", 1_000_000, 4), ("c4_clustered", "clustered", 1_000_000, 4)])
def test_pruned_full_size_digests(name, kind, n, cfg, cuda_lib, device):
    """WHOLE output tensors against the SHA-256 digests of the compiled reference (the same digests the brute-force path
    is held to, tests/golden/digests_full.json)."""
    meta = FULL_DIGESTS[name]
    x1, x2 = synthetic.make_pair(kind, meta["b"], n, n, cfg)
    assert sha(x1.numpy()) == meta["input_sha256"]["xyz1"] and sha(x2.numpy()) == meta["input_sha256"]["xyz2"]
    got = pruned(x1, x2, device)
    for k, v in zip(("dist1", "dist2", "idx1", "idx2"), got):
        assert sha(v) == meta["sha256"][k], f"{name}: pruned {k} differs from the reference's full-size output"


def test_pruned_c2_batch_digest(cuda_lib, device):
    meta = FULL_DIGESTS["c2_uniform"]
    x1, x2 = synthetic.make_pair("uniform", meta["b"], meta["n"], meta["m"], 2)
    assert sha(x1.numpy()) == meta["input_sha256"]["xyz1"]
    got = pruned(x1, x2, device)
    for k, v in zip(("dist1", "dist2", "idx1", "idx2"), got):
        assert sha(v) == meta["sha256"][k], f"c2: pruned {k} differs from the reference's full-size output"


def test_pruned_ten_million_points_sampled_rows(cuda_lib, device):
    """N = M = 10^7 (the brute-force scan would take ~25 s per forward): three box levels.  160 sampled rows per direction
    against the oracle, plus the size-independent properties: the reported neighbour is at the reported distance, and a
    second call on the swapped clouds returns the swapped results."""
    n = 10_000_000
    g = torch.Generator().manual_seed(4242)
    x1 = torch.rand(1, n, 3, generator=g)
    x2 = torch.rand(1, n, 3, generator=g)
    x2[0, 5_000_000:5_000_100] = x1[0, 100:200]  # exact copies: distance 0 at a known index
    a, c = x1.to(device), x2.to(device)
    d1, d2, i1, i2 = F.nnd_forward_pruned_raw(a, c)
    e2, e1, j2, j1 = F.nnd_forward_pruned_raw(c, a)
    torch.cuda.synchronize()
    assert torch.equal(d1, e1) and torch.equal(i1, j1) and torch.equal(d2, e2) and torch.equal(i2, j2)
    assert float(d2[0, 5_000_000:5_000_100].abs().max()) == 0.0 and torch.equal(i2[0, 5_000_000:5_000_100].cpu(),
                                                                                torch.arange(100, 200, dtype=torch.int32))
    rows = torch.randperm(n, generator=g)[:160].sort().values
    w1 = oracle.query(x1[:, rows].numpy(), x2.numpy(), threads=16)
    w2 = oracle.query(x2[:, rows].numpy(), x1.numpy(), threads=16)
    assert d1[:, rows].cpu().numpy().tobytes() == w1[0].tobytes() and i1[:, rows].cpu().numpy().tobytes() == w1[1].tobytes()
    assert d2[:, rows].cpu().numpy().tobytes() == w2[0].tobytes() and i2[:, rows].cpu().numpy().tobytes() == w2[1].tobytes()
    nb = c[0, i1[0].long()]  # the reported neighbour is at the reported distance (same rounding sequence in torch fp32)
    dx = nb - a[0]
    assert torch.equal((dx[:, 0] * dx[:, 0] + dx[:, 1] * dx[:, 1]) + dx[:, 2] * dx[:, 2], d1[0])
