"""Dev tool: the single-launch small-cloud kernels (csrc/nnd_small.cuh) -- device time per forward + backward at the
shapes the reference's callers use, every configuration of the forward, the crossover against the general path, and
the host cost of the public API around them.

    python tools/small_time.py [configs] [crossover] [api]

Device times are CUDA-event times over a CUDA graph of REPS forward+backward pairs (no launch gaps from Python), so
they are what the GPU spends; "eager" lines add the per-call host path of functional.nnd_forward_raw/backward_raw.
"""
import sys
import time
from pathlib import Path

import torch

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from pointcloudregistration_b200 import _cabi  # noqa: E402
from pointcloudregistration_b200 import functional as F  # noqa: E402
from pointcloudregistration_b200 import synthetic  # noqa: E402

dev = torch.device("cuda:0")
lib = _cabi.load()
PEAK = 148 * 128 * 1.965e9
REPS = 20


def setup(b, n, m):
    x1, x2 = synthetic.make_pair("uniform", b, n, m, 4)
    a, c = x1.to(dev), x2.to(dev)
    gd1, gd2 = (t.to(dev) for t in synthetic.grad_pair(b, n, m, 4, mean=True))
    d1, d2 = torch.empty(b, n, device=dev), torch.empty(b, m, device=dev)
    i1 = torch.empty(b, n, dtype=torch.int32, device=dev)
    i2 = torch.empty(b, m, dtype=torch.int32, device=dev)
    g1, g2 = torch.empty_like(a), torch.empty_like(c)
    return a, c, gd1, gd2, d1, d2, i1, i2, g1, g2


def graph_time(b, n, m, fwd=True, bwd=True):
    """ms per (forward [+ backward]) on the device, from a graph of REPS repetitions."""
    a, c, gd1, gd2, d1, d2, i1, i2, g1, g2 = setup(b, n, m)
    nbytes = lib.nnd_workspace_bytes(b, n, m)
    ws = torch.empty(max(nbytes, 256), dtype=torch.uint8, device=dev)

    def body():
        st = torch.cuda.current_stream().cuda_stream
        if fwd:
            rc = lib.nnd_forward_cuda_f32(a.data_ptr(), c.data_ptr(), b, n, m, d1.data_ptr(), d2.data_ptr(), i1.data_ptr(),
                                          i2.data_ptr(), ws.data_ptr(), ws.numel(), st)
            assert rc == 0, rc
        if bwd:
            rc = lib.nnd_backward_cuda_f32(a.data_ptr(), c.data_ptr(), gd1.data_ptr(), gd2.data_ptr(), i1.data_ptr(),
                                           i2.data_ptr(), b, n, m, g1.data_ptr(), g2.data_ptr(), st)
            assert rc == 0, rc

    s = torch.cuda.Stream()
    with torch.cuda.stream(s):
        body()
        body()
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        for _ in range(REPS):
            body()
    g.replay()
    torch.cuda.synchronize()
    best = 1e9
    for _ in range(5):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        g.replay()
        e1.record()
        torch.cuda.synchronize()
        best = min(best, e0.elapsed_time(e1) / REPS)
    return best


def eager_time(b, n, m):
    a, c, gd1, gd2, *_ = setup(b, n, m)
    out = F.nnd_forward_raw(a, c)

    def step():
        o = F.nnd_forward_raw(a, c)
        F.nnd_backward_raw(a, c, gd1, gd2, o[2], o[3])

    for _ in range(5):
        step()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(200):
        step()
    torch.cuda.synchronize()
    return (time.perf_counter() - t0) / 200 * 1e3, out


def line(tag, b, n, m, ms_f, ms_fb):
    pairs = 2.0 * b * n * m
    print(f"{tag:34s} B={b:4d} N={n:6d} M={m:6d}  fwd {ms_f * 1e3:8.2f} us  fwd+bwd {ms_fb * 1e3:8.2f} us  "
          f"{pairs / (ms_fb * 1e-3) / 1e9:8.1f} Gpairs/s  fwd = {pairs * 8 / (ms_f * 1e-3) / PEAK * 100:5.1f}% of FP32 issue peak",
          flush=True)


SHAPES = [(1, 2048, 2048), (256, 256, 256), (1, 3000, 2900), (64, 256, 256), (16, 2048, 1024), (1, 4096, 4096)]


def main():
    what = set(sys.argv[1:]) or {"configs", "crossover", "api"}
    print("== default configuration, small path vs general path (device time, graph of 20 repetitions)")
    for b, n, m in SHAPES:
        line("small (auto config)", b, n, m, graph_time(b, n, m, bwd=False), graph_time(b, n, m))
        before = lib.nnd_set_small_max_points(0)
        line("general (prepare+scan+finalize)", b, n, m, graph_time(b, n, m, bwd=False), graph_time(b, n, m))
        lib.nnd_set_small_max_points(before)
    if "configs" in what:
        print("== every forward configuration (warps, queries/lane, target split), forward only")
        for b, n, m in SHAPES[:4] + [(32, 1024, 1024), (8, 4096, 4096)]:
            res = []
            for w in (2, 4, 8, 16):
                for q in (1, 2, 4, 8):
                    for t in (1, 2, 4, 8, 16):
                        if w % t:
                            continue
                        lib.nnd_set_small_config(w, q, t)
                        res.append((graph_time(b, n, m, bwd=False), (w, q, t)))
            lib.nnd_set_small_config(0, 0, 0)
            auto = graph_time(b, n, m, bwd=False)
            res.sort()
            print(f"B={b} N={n} M={m}: auto {auto * 1e3:.2f} us; best " +
                  ", ".join(f"{c}: {t_ * 1e3:.2f} us" for t_, c in res[:10]), flush=True)
    if "crossover" in what:
        print("== crossover: single-launch vs general path, B=1, N=M (forward+backward)")
        before = lib.nnd_small_max_points()
        for n in (512, 1024, 2048, 4096, 6144, 8192, 12288, 16384):
            lib.nnd_set_small_max_points(16384)
            ts_f, ts = graph_time(1, n, n, bwd=False), (graph_time(1, n, n) if 2 * n * 12 <= 200 * 1024 else float("nan"))
            lib.nnd_set_small_max_points(0)
            tg_f, tg = graph_time(1, n, n, bwd=False), graph_time(1, n, n)
            print(f"N=M={n:6d}: small fwd {ts_f * 1e3:8.2f} us fwd+bwd {ts * 1e3:8.2f} us | general fwd {tg_f * 1e3:8.2f} us "
                  f"fwd+bwd {tg * 1e3:8.2f} us", flush=True)
        for b in (8, 32):
            for n in (2048, 4096, 8192):
                lib.nnd_set_small_max_points(16384)
                ts_f = graph_time(b, n, n, bwd=False)
                lib.nnd_set_small_max_points(0)
                tg_f = graph_time(b, n, n, bwd=False)
                print(f"B={b:3d} N=M={n:6d}: small fwd {ts_f * 1e3:8.2f} us | general fwd {tg_f * 1e3:8.2f} us", flush=True)
        lib.nnd_set_small_max_points(before)
    if "api" in what:
        print("== public API, eager (host path included): functional.nnd_forward_raw + nnd_backward_raw")
        for b, n, m in SHAPES[:3]:
            ms, _ = eager_time(b, n, m)
            print(f"B={b} N={n} M={m}: {ms * 1e3:.1f} us per forward+backward (eager, back to back)", flush=True)
        import cProfile
        import pstats

        a, c, gd1, gd2, *_ = setup(1, 2048, 2048)
        import torch_nndistance as NND

        def autograd_step():
            p = a.detach().requires_grad_(True)
            q = c.detach().requires_grad_(True)
            d1, d2 = NND.nnd(p, q)
            (d1.mean() + d2.mean()).backward()

        for _ in range(10):
            autograd_step()
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        for _ in range(200):
            autograd_step()
        torch.cuda.synchronize()
        print(f"C1 through NND.nnd + autograd: {(time.perf_counter() - t0) / 200 * 1e6:.1f} us per step")
        pr = cProfile.Profile()
        pr.enable()
        for _ in range(200):
            autograd_step()
        torch.cuda.synchronize()
        pr.disable()
        pstats.Stats(pr).sort_stats("cumulative").print_stats(18)


if __name__ == "__main__":
    main()
