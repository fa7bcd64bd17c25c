"""GPU tests of the graph-captured chamfer iteration (pointcloudregistration_b200/ndp_loop.py) against the eager
pytorch3d-shaped loss + autograd (the path registration.py:236-261 takes)."""
import sys
from pathlib import Path

import pytest
import torch

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))

from pointcloudregistration_b200 import ndp_loop, synthetic  # noqa: E402
from pointcloudregistration_b200 import pytorch3d_compat as P3D  # noqa: E402

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("trunc", [1e9, 0.0005])
@pytest.mark.parametrize("use_graph", [True, False])
def test_loop_matches_eager_over_iterations(cuda_lib, device, trunc, use_graph):
    n, m = 3000, 2900
    x0, y = synthetic.make_pair("clustered", 1, n, m, 95)
    y = y.to(device)
    loop = ndp_loop.ChamferLoop(1, n, m, device, trunc=trunc, use_graph=use_graph)
    loop.set_target(y)
    xa = x0.to(device).clone()
    xb = x0.to(device).clone()
    for it in range(6):
        loss, g = loop.step(xa)
        xe = xb.clone().requires_grad_(True)
        want = P3D.compute_truncated_chamfer_distance(xe, y, trunc=trunc)
        want.backward()
        assert torch.allclose(loss, want.detach(), rtol=3e-6, atol=0), (it, float(loss), float(want))
        scale = xe.grad.abs().max()
        assert (g - xe.grad).abs().max() <= 1e-5 * scale
        xa = xa - 0.01 * g          # registration.py:259-261 (stand-in for the optimiser step)
        xb = xb - 0.01 * xe.grad
    assert loop.replays == 6
    if trunc < 1:
        assert 0 < int((loop.grad_x.abs().sum(-1) == 0).sum())  # truncated rows get no gradient


def test_autograd_function_trains_a_warp(cuda_lib, device):
    n, m = 1024, 1500
    x0, y = synthetic.make_pair("uniform", 1, n, m, 96)
    loop = ndp_loop.ChamferLoop(1, n, m, device)
    loop.set_target(y.to(device))
    shift = torch.zeros(3, device=device, requires_grad=True)  # a 3-parameter "warp": x + shift
    opt = torch.optim.Adam([shift], lr=0.01)
    first = None
    for it in range(30):
        opt.zero_grad()
        loss = ndp_loop.ChamferLoopFunction.apply(x0.to(device) + 0.2 + shift, loop)
        loss.backward()
        opt.step()
        if it % 10 == 0:  # the early-stop read, every k-th iteration only
            v = loop.loss_value()
            first = v if first is None else first
    assert loop.loss_value() < first and shift.grad is not None


@pytest.mark.parametrize("b,n,m", [(1, 3000, 2900), (3, 5, 0), (2, 300000, 170001), (1, 0, 0)])
def test_fused_reduce_matches_float64_numpy(cuda_lib, device, b, n, m):
    import numpy as np

    from pointcloudregistration_b200 import functional as F

    g = torch.Generator().manual_seed(97)
    d1, d2 = torch.rand(b, n, generator=g), torch.rand(b, m, generator=g)
    if n:
        d1[0, 0] = 0.25          # exactly the truncation value: dropped (>=)
        d1[0, n // 2] = float("nan")  # NaN is not >= trunc: kept, as in loss.py:151-154
    trunc, s1, s2 = 0.25, 1.0 / max(b * n, 1), 0.5
    ws = F.chamfer_reduce_workspace(device)
    ws.fill_(0xa5)  # the scratch carries no state: garbage in it (or a ticket left by an aborted launch) must not matter
    loss = torch.full((), 7.0, device=device)
    gd1, gd2 = torch.empty(b, n, device=device), torch.empty(b, m, device=device)
    for rep in range(3):
        F.chamfer_reduce_raw(d1.to(device), d2.to(device), trunc, s1, s2, loss, gd1, gd2, ws)
        a1, a2 = d1.double().numpy(), d2.double().numpy()
        k1, k2 = ~(a1 >= trunc), ~(a2 >= trunc)
        want = s1 * a1[k1].sum() + s2 * a2[k2].sum()
        got = float(loss)
        if np.isnan(want):
            assert np.isnan(got)
        else:
            assert abs(got - want) <= 2e-7 * abs(want) + 1e-30
        assert np.array_equal(gd1.cpu().numpy(), np.where(k1, np.float32(s1), np.float32(0)))
        assert np.array_equal(gd2.cpu().numpy(), np.where(k2, np.float32(s2), np.float32(0)))
        d1 = d1.nan_to_num(0.1)  # second and third rounds: finite
    ws.fill_(0xff)
    F.chamfer_reduce_raw(d1.to(device), d2.to(device), trunc, s1, s2, loss, None, None, ws)  # no gradient wanted


class _Warp(torch.nn.Module):
    """A small stand-in for one NDP pyramid level (nets.py:66-143 is out of scope): x + MLP(x)."""

    def __init__(self, width=64):
        super().__init__()
        self.net = torch.nn.Sequential(torch.nn.Linear(3, width), torch.nn.ReLU(), torch.nn.Linear(width, width),
                                       torch.nn.ReLU(), torch.nn.Linear(width, 3))
        with torch.no_grad():
            self.net[-1].weight.mul_(0.01)

    def forward(self, x):
        return x + self.net(x)


def _iteration(n, m, seed, device, use_graph, exact=True):
    torch.manual_seed(seed)
    warp = _Warp().to(device)
    opt = torch.optim.Adam(warp.parameters(), lr=1e-3, capturable=True)
    x0, y = synthetic.make_pair("clustered", 1, n, m, seed)
    return ndp_loop.GraphedIteration(warp, opt, x0[0].to(device) + 0.05, y[0].to(device), trunc=1e9,
                                     use_graph=use_graph, exact=exact), warp


@pytest.mark.parametrize("n,m", [(3000, 2900), (100_000, 100_000)])
def test_whole_iteration_graph_matches_the_eager_loop_for_500_iterations(cuda_lib, device, n, m):
    """warp MLP forward -> chamfer -> backward -> Adam step captured as ONE CUDA graph (registration.py:231-261) against
    the same sequence run eagerly, 500 iterations, deterministic backward on both sides: the loss curve (read every
    10th iteration, the early-stop protocol) agrees to fp32 rounding and so do the final warped points."""
    ga, wa = _iteration(n, m, 98, device, use_graph=True)
    gb, wb = _iteration(n, m, 98, device, use_graph=False)
    for pa, pb in zip(wa.parameters(), wb.parameters()):
        assert torch.equal(pa, pb)
    ha = ga.run(500, read_every=10, loss_tol=0.0, break_threshold_ratio=0.0)
    hb = gb.run(500, read_every=10, loss_tol=0.0, break_threshold_ratio=0.0)
    assert len(ha) == len(hb) == 51 and ga.iterations == 500
    for (ia, la), (ib, lb) in zip(ha, hb):
        assert ia == ib and abs(la - lb) <= 2e-6 * abs(lb), (ia, la, lb)
    assert ha[-1][1] < 0.9 * ha[0][1]  # it trains
    assert torch.allclose(ga.warped, gb.warped, rtol=0, atol=2e-6)
    for pa, pb in zip(wa.parameters(), wb.parameters()):
        assert torch.allclose(pa, pb, rtol=1e-4, atol=1e-6)


def test_whole_iteration_early_stop_protocol(cuda_lib, device):
    """run() applies registration.py:250-256's rules on the losses it reads: stops once the loss is below the
    tolerance, and the atomic backward (default) trains too."""
    g, _ = _iteration(2000, 2000, 99, device, use_graph=True, exact=False)
    hist = g.run(300, read_every=10, loss_tol=1e9)  # any loss is below 1e9: stops at the first read
    assert len(hist) == 1 and g.iterations == 1
    hist = g.run(200, read_every=20, loss_tol=0.0, break_threshold_ratio=0.0)
    assert [it for it, _ in hist] == list(range(0, 200, 20)) + [199] and hist[-1][1] < hist[0][1]


@pytest.mark.parametrize("use_graph", [True, False])
def test_loop_with_the_pruned_search_is_bit_identical_to_the_brute_force_loop(cuda_lib, device, use_graph):
    """ChamferLoop(search='pruned'): the exact pruned forward (csrc/nnd_pruned.cuh, incl. the library sort) captured
    into the CUDA graph.  Same bits out of the forward, so -- with the deterministic-order epilogue and the same
    backward kernel on the same inputs -- the same loss bits, iteration after iteration."""
    n, m = 30000, 26000
    x0, y = synthetic.make_pair("crosscopy", 1, n, m, 97)
    y = y.to(device)
    a = ndp_loop.ChamferLoop(1, n, m, device, use_graph=use_graph, search="pruned")
    b = ndp_loop.ChamferLoop(1, n, m, device, use_graph=use_graph, search="brute")
    assert a.pruned and not b.pruned
    a.set_target(y)
    b.set_target(y)
    xa = x0.to(device).clone()
    for it in range(8):
        la, ga = a.step(xa)
        lb, gb = b.step(xa)
        assert torch.equal(a.idx1, b.idx1) and torch.equal(a.idx2, b.idx2), it
        assert torch.equal(a.dist1, b.dist1) and torch.equal(a.dist2, b.dist2), it
        assert torch.equal(la, lb), (it, float(la), float(lb))
        scale = gb.abs().max()
        assert (ga - gb).abs().max() <= 1e-5 * scale  # atomic summation order
        xa = xa - 5.0 * gb
    assert a.replays == 8


@pytest.mark.parametrize("use_graph", [True, False])
def test_pruned_loop_reuses_the_target_tiles_and_a_stale_source_order(cuda_lib, device, use_graph):
    """nnd_forward_pruned_reuse_cuda_f32 through ChamferLoop(search='pruned', resort_every=4): the target is prepared
    once per set_target(), the moving cloud keeps its curve order for 3 of 4 steps.  Even with the cloud moving far
    (every step, plus a jump) every iteration returns the brute-force loop's bits; a new target resets the reuse."""
    n, m = 20000, 23000
    x0, y = synthetic.make_pair("clustered", 1, n, m, 98)
    _, y2 = synthetic.make_pair("uniform", 1, n, m, 99)
    y, y2 = y.to(device), y2.to(device)
    a = ndp_loop.ChamferLoop(1, n, m, device, use_graph=use_graph, search="pruned", resort_every=4)
    b = ndp_loop.ChamferLoop(1, n, m, device, use_graph=use_graph, search="brute")
    a.set_target(y)
    b.set_target(y)
    xa = x0.to(device).clone()
    seen = []
    for it in range(14):
        if it == 9:  # a new target in the middle of the loop
            a.set_target(y2)
            b.set_target(y2)
        la, ga = a.step(xa)
        lb, gb = b.step(xa)
        seen.append(a._since_sort)
        assert torch.equal(a.idx1, b.idx1) and torch.equal(a.idx2, b.idx2), it
        assert torch.equal(a.dist1, b.dist1) and torch.equal(a.dist2, b.dist2), it
        assert torch.equal(la, lb), (it, float(la), float(lb))
        xa = xa - 20.0 * gb
        if it == 5:  # a jump the stale order knows nothing about (the order is only a pruning aid)
            xa = xa.flip(1) * 0.5 + 0.3
    assert seen == [0, 1, 2, 3, 0, 1, 2, 3, 0, 0, 1, 2, 3, 0]  # steps since the moving cloud was last sorted
    if use_graph:
        assert sorted(a._graphs) == [0, 2, 3]
