"""CPU property test of the inequality the exact pruned search rests on (csrc/nnd_pruned.cuh):

    for every target t inside a box and every query q (inside a query box):
        pr_safe(box distance evaluated in fp32)  <=  fl-distance(q, t)

where fl-distance is the reference's separately rounded fp32 sequence (my_lib.cpp:12-15; the oracle evaluates it) and
pr_safe shrinks the bound by 2^-20 relative and 2^-100 absolute.  A box is skipped only when its bound EXCEEDS the current
best, so this inequality is what makes skipping exact.  numpy float32 restates both sides; the GPU may contract the
bound's multiply-adds into FMAs, so the bound is also evaluated in float64 and inflated by 2^-21 (more than any fp32
evaluation order can add) before it is shrunk -- the inequality must still hold.  No GPU needed."""
import numpy as np
import pytest

F = np.float32
REL = F(1.0) - F(9.5367431640625e-7)      # pr_safe: 1 - 2^-20
ABS = F(7.8886090522101181e-31)           # 2^-100


def fl_dist(q, t):
    """my_lib.cpp:12-15 in numpy float32: target - query, (xx + yy) + zz, every operation rounded separately."""
    d = (t - q).astype(F)
    return ((d[..., 0] * d[..., 0]).astype(F) + (d[..., 1] * d[..., 1]).astype(F)).astype(F) + (d[..., 2] * d[..., 2]).astype(F)


def safe(lb):
    return (lb.astype(F) * REL).astype(F) - ABS


def gap32(lo_a, hi_a, lo_b, hi_b):
    """per-axis gap between two intervals, fp32 (pr_box_box / pr_point_box with a degenerate box)"""
    return np.maximum(np.maximum((lo_a - hi_b).astype(F), (lo_b - hi_a).astype(F)), F(0))


def bound32(lo_a, hi_a, lo_b, hi_b):
    g = gap32(lo_a, hi_a, lo_b, hi_b)
    return ((g[..., 0] * g[..., 0]).astype(F) + (g[..., 1] * g[..., 1]).astype(F)).astype(F) + (g[..., 2] * g[..., 2]).astype(F)


def bound64_inflated(lo_a, hi_a, lo_b, hi_b):
    g = np.maximum(np.maximum(lo_a.astype(np.float64) - hi_b, lo_b.astype(np.float64) - hi_a), 0.0)
    lb = (g * g).sum(-1) * (1.0 + 2.0 ** -21)
    with np.errstate(over="ignore"):
        return np.nextafter(lb.astype(F), F(np.inf))  # rounded UP to fp32: the largest any fp32 evaluation could give


@pytest.mark.parametrize("scale", [1e-22, 1e-12, 1e-3, 1.0, 37.5, 4096.0, 1e9, 1e18])
@pytest.mark.parametrize("spread", [0.0, 1e-6, 1e-2, 1.0])
def test_shrunk_box_bounds_never_exceed_the_rounded_distance(scale, spread):
    rng = np.random.default_rng(int(abs(np.log10(scale)) * 100) + int(spread * 1000) + 7)
    n = 4000
    with np.errstate(over="ignore", invalid="ignore", under="ignore"):
        centre_t = (rng.normal(size=(n, 1, 3)) * scale).astype(F)
        centre_q = (centre_t[:, 0] + rng.normal(size=(n, 3)) * scale * rng.choice([1e-7, 1e-3, 1.0, 50.0], size=(n, 1))).astype(F)
        # a "tile" of 16 targets around centre_t and a "warp" of 8 queries around centre_q, sizes from tiny to large
        t = (centre_t + rng.normal(size=(n, 16, 3)) * scale * spread).astype(F)
        q = (centre_q[:, None] + rng.normal(size=(n, 8, 3)) * scale * spread * 0.5).astype(F)
        tlo, thi = t.min(1), t.max(1)
        qlo, qhi = q.min(1), q.max(1)
        d = fl_dist(q[:, :, None, :], t[:, None, :, :])          # [n, 8, 16] every query against every target of its tile
        dmin_box = d.min((1, 2))                                  # the smallest rounded distance between the two boxes' points
        dmin_pt = d.min(2)                                        # per query
        for bound in (bound32, bound64_inflated):
            bb = safe(bound(qlo, qhi, tlo, thi))                  # pr_box_box
            ok = ~(bb > dmin_box)                                 # the kernel prunes on `bound > best`: NaN never prunes
            assert ok.all(), (scale, spread, bound.__name__, float(bb[~ok][0]), float(dmin_box[~ok][0]))
            pb = safe(bound(q, q, tlo[:, None, :], thi[:, None, :]))  # pr_point_box: the query as a degenerate box
            ok = ~(pb > dmin_pt)
            assert ok.all(), (scale, spread, bound.__name__, float(pb[~ok][0]), float(dmin_pt[~ok][0]))


def test_bounds_with_nonfinite_boxes_never_prune_a_finite_winner():
    """Boxes are built with NaN-dropping min / max; +-inf coordinates stay in them.  Whatever comes out of the bound then
    (inf, NaN from inf - inf), `bound > best` must be False whenever the box holds a target at a finite distance."""
    rng = np.random.default_rng(3)
    with np.errstate(all="ignore"):
        for _ in range(200):
            t = rng.random((16, 3)).astype(F)
            q = rng.random((8, 3)).astype(F)
            k = int(rng.integers(1, 4))
            t[rng.integers(0, 16, k), rng.integers(0, 3, k)] = rng.choice(np.array([np.inf, -np.inf, 3e19], F), k)
            if rng.random() < 0.5:
                q[rng.integers(0, 8), rng.integers(0, 3)] = rng.choice(np.array([np.inf, -np.inf], F))
            tlo, thi = np.fmin.reduce(t, 0), np.fmax.reduce(t, 0)
            qlo, qhi = np.fmin.reduce(q, 0), np.fmax.reduce(q, 0)
            d = fl_dist(q[:, None, :], t[None, :, :])
            finite = np.isfinite(d)
            if not finite.any():
                continue
            best = np.where(finite, d, np.inf).min()
            g = np.fmax(np.fmax((qlo - thi).astype(F), (tlo - qhi).astype(F)), F(0))  # fmaxf drops NaN gaps
            lb = safe(((g[0] * g[0]).astype(F) + (g[1] * g[1]).astype(F)).astype(F) + (g[2] * g[2]).astype(F))
            assert not (lb > best), (float(lb), float(best))
