#!/usr/bin/env python
"""bench.py -- chamfer / nnd forward+backward throughput on B200 (BASELINE.json metric).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--workload c4] [--kind uniform]
    python bench.py --impl reference ...      # the reference's own CPU nnd on the host cores
    torchrun --nnodes=1 --nproc-per-node N ... bench.py --gpus N ...

One "step" = one nnd forward (dist1, dist2, idx1, idx2) + one backward (gradxyz1, gradxyz2) over one
pair of synthetic clouds.  Pairs per step = 2*B*N*M (both directions).  Prints ONE JSON line.

  value      Gpairs/s, whole job, inputs resident in HBM when the timed region starts
  e2e        the same metric through the public API (torch_nndistance.nnd + autograd) with HOST
             (pinned) inputs: H2D of both clouds, D2H of both gradients + the loss, every step
  roofline   the dominant kernel (the scan: nnd_scan_symc_kernel at C4; roofline.kernel names it): 8
             separately-rounded FP32 ops per evaluated distance (3 sub, 3 mul, 2 add -- FMA is forbidden
             by bit-exactness) / the scan's CUDA-event duration measured inside the timed steps, against
             the FP32 issue peak 148 SM x 128 lanes x sm_max_mhz
  cpu_baseline  the reference's own my_lib.cpp (oracle/_ref, unmodified) on all host cores, on a
             bounded query sample of the same workload, same run (N=1, rank 0 only)

The oracle / reference CPU code is executed ONLY in the cpu_baseline leg and in --impl reference.
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import subprocess
import sys
import threading
import time
from pathlib import Path

ROOT = Path(__file__).resolve().parent
if str(ROOT) not in sys.path:
    sys.path.insert(0, str(ROOT))

METRIC = "chamfer NN Gpairs/s fwd+bwd"
UNIT = "Gpairs/s"
FP32_OPS_PER_PAIR = 8  # 3 sub + 3 mul + 2 add, each separately rounded
BWD_BYTES_PER_ROW = 56
SMS, LANES = 148, 128

WORKLOADS = {  # BASELINE.json configs
    "c1": dict(b=1, n=2048, m=2048, config=1, kind="uniform", desc="C1 B=1 N=M=2048 (DIP patch size)"),
    "c2": dict(b=32, n=16384, m=16384, config=2, kind="uniform", desc="C2 DIP training-shape batch B=32 N=M=16384"),
    "c3": dict(b=1, n=100_000, m=100_000, config=3, kind="uniform", desc="C3 NDP/C2P-Net pair N=M=100k"),
    "c4": dict(b=1, n=1_000_000, m=1_000_000, config=4, kind="uniform", desc="C4 large-cloud chamfer N=M=1M"),
    "c5": dict(b=1, n=262_144, m=262_144, config=5, kind="lattice", desc="C5 tie stress N=M=262144"),
}


_RESULT_FD = None


def claim_stdout():
    """The contract is ONE JSON line on stdout, but libraries write there too (NCCL prints its version line to stdout
    when it initialises).  From here on everything written to fd 1 lands on stderr; emit() owns the real stdout."""
    global _RESULT_FD
    if _RESULT_FD is None:
        sys.stdout.flush()
        _RESULT_FD = os.dup(1)
        os.dup2(2, 1)


def emit(line: dict) -> None:
    data = (json.dumps(line) + "\n").encode()
    if _RESULT_FD is None:
        sys.stdout.write(data.decode())
        sys.stdout.flush()
    else:
        os.write(_RESULT_FD, data)


def load_peaks():
    p = ROOT / "MEASURED_PEAKS.json"
    if p.exists():
        d = json.loads(p.read_text())
        return float(d.get("hbm_gbs", 6650.0)), float(d.get("sm_max_mhz", 1965.0)), "measured (MEASURED_PEAKS.json)"
    return 6650.0, 1965.0, "fallback (B200_PROFILING.md)"


# --------------------------------------------------------------------------------------------- clocks
class ClockSampler:
    """Samples SM clock + throttle reasons of one GPU every 100 ms while the timed region runs."""

    BAD = {"hw_slowdown": 0x8, "sw_thermal_slowdown": 0x20, "hw_thermal_slowdown": 0x40, "hw_power_brake": 0x80}
    NOTE = {"sw_power_cap": 0x4}

    def __init__(self, index: int):
        self.index, self.samples, self.reasons = index, [], set()
        self.max_mhz, self._stop, self._thr = None, threading.Event(), None
        try:
            import pynvml

            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
        except Exception:  # noqa: BLE001
            self.nv = None

    def _loop(self):
        while not self._stop.is_set():
            try:
                self.samples.append(self.nv.nvmlDeviceGetClockInfo(self.h, self.nv.NVML_CLOCK_SM))
                mask = self.nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                for name, bit in {**self.BAD, **self.NOTE}.items():
                    if mask & bit:
                        self.reasons.add(name)
            except Exception:  # noqa: BLE001
                pass
            self._stop.wait(0.1)

    def __enter__(self):
        if self.nv is not None:
            self._thr = threading.Thread(target=self._loop, daemon=True)
            self._thr.start()
        return self

    def __exit__(self, *exc):
        self._stop.set()
        if self._thr is not None:
            self._thr.join()

    def summary(self):
        if not self.samples:
            return {"sm_mhz": None, "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons), "samples": 0}
        return {"sm_mhz": statistics.median(self.samples), "sm_max_mhz": self.max_mhz,
                "reasons": sorted(self.reasons), "samples": len(self.samples)}


# --------------------------------------------------------------------------------------------- CPU arm
def cpu_arm(w, steps: int, warmup: int, seconds_per_step: float, kind: str):
    """Times the reference's CPU nnd (oracle/_ref, else the oracle port) on all host cores.

    Every core runs the reference's own nnd_forward + nnd_backward on a disjoint slice of cloud-1
    query rows against the full cloud 2 (both directions, exactly what nnd_forward does for an
    [b, q, 3] x [b, m, 3] call); slices are sized for ~seconds_per_step.  Workers are separate
    processes (the reference's pybind entry points hold the GIL) released together by a start time.
    """
    cores = len(os.sched_getaffinity(0))
    if getattr(cpu_arm, "max_cores", None):
        cores = min(cores, cpu_arm.max_cores)
    b, n, m = w["b"], w["n"], w["m"]
    per_core_rate = 0.30e9  # pairs/s/core; only used to size the sample
    q = int(seconds_per_step * per_core_rate / (2.0 * b * m))
    q = max(1, min(q, n // cores if n >= cores else 1))
    workers = min(cores, n)
    start_at = time.time() + 25.0  # python + torch import of every worker happens before this
    procs = []
    for i in range(workers):
        cmd = [sys.executable, str(ROOT / "oracle" / "cpu_worker.py"), "--kind", kind, "--b", str(b), "--n", str(n),
               "--m", str(m), "--cloud", w["kind"], "--config", str(w["config"]), "--lo", str(i * q),
               "--hi", str((i + 1) * q), "--steps", str(steps), "--warmup", str(warmup), "--start-at", str(start_at)]
        procs.append(subprocess.Popen(cmd, stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True))
    per_worker = []
    impl_kind = None
    for p in procs:
        out, err = p.communicate()
        if p.returncode != 0:
            raise RuntimeError(f"cpu worker failed: {err[-2000:]}")
        rec = json.loads(out.strip().splitlines()[-1])
        per_worker.append(rec["step_seconds"])
        impl_kind = rec["kind"]
    # step time = slowest worker of that step (they run concurrently on all cores)
    step_s = [max(wk[i] for wk in per_worker) for i in range(steps)]
    pairs_per_step = 2.0 * b * (q * workers) * m
    total_s = sum(step_s)
    gpairs = pairs_per_step * steps / total_s / 1e9
    sample = (f"{workers} workers x {q} cloud-1 query rows each (of N={n}) against the full M={m} cloud, both "
              f"directions + backward = {pairs_per_step:.3e} pairs per step, {steps} step(s); rate is per pair, "
              f"so it extrapolates linearly to the full workload")
    return dict(value=gpairs, unit=UNIT, cores=workers, kind=impl_kind, sample=sample), total_s / steps * 1e3


def run_reference_arm(args, w):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    steps, warmup = args.steps, args.warmup
    kind = "reference" if (ROOT / "oracle" / "_ref" / "nnd_ref_cpu.so").exists() else "port"
    base, ms = cpu_arm(w, steps, warmup, args.cpu_seconds_per_step, kind)
    line = {
        "impl": "reference", "metric": METRIC, "value": base["value"], "unit": UNIT, "n_gpus": args.gpus,
        "steps": steps, "warmup": warmup, "ms_per_step": ms, "higher_is_better": True, "scaling": "strong",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": w["desc"], "cloud": w["kind"], "b": w["b"], "n": w["n"], "m": w["m"],
                   "note": "each step is a bounded query sample of the workload on the host cores"},
        "cpu_baseline": base,
        "e2e": {"value": base["value"], "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    emit(line)


# --------------------------------------------------------------------------------------------- reference GPU kernel
def run_reference_cuda_arm(args, w):
    """Comparison row: the reference's OWN CUDA kernels (nnd_cuda.cu, unmodified, recompiled for
    sm_100a into oracle/_ref/nnd_ref_cuda.so) driven the way torch_nndistance/__init__.py:11-57 drives them, same
    metric, same inputs, 1 GPU.  Not a parity oracle (FMA-contracted); the line reports how far its outputs are from
    the B200 path's (which is bit-exact against the reference CPU code)."""
    import importlib.util

    import torch

    from pointcloudregistration_b200 import functional as F
    from pointcloudregistration_b200 import synthetic

    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    so = ROOT / "oracle" / "_ref" / "nnd_ref_cuda.so"
    if not so.exists():
        emit({"impl": "reference_cuda", "unavailable": "oracle/_ref/nnd_ref_cuda.so not built"})
        return
    spec = importlib.util.spec_from_file_location("nnd_ref_cuda", str(so))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    dev = torch.device("cuda", 0)
    torch.cuda.set_device(dev)
    b, n, m = w["b"], w["n"], w["m"]
    x1h, x2h = synthetic.make_pair(w["kind"], b, n, m, w["config"])
    gd1h, gd2h = synthetic.grad_pair(b, n, m, w["config"], mean=True)
    x1, x2, gd1, gd2 = x1h.to(dev), x2h.to(dev), gd1h.to(dev), gd2h.to(dev)

    def step():
        d1 = torch.zeros(b, n, device=dev)
        d2 = torch.zeros(b, m, device=dev)
        i1 = torch.zeros(b, n, dtype=torch.int32, device=dev)
        i2 = torch.zeros(b, m, dtype=torch.int32, device=dev)
        mod.nnd_forward_cuda(x1, x2, d1, d2, i1, i2)  # __init__.py:29
        g1, g2 = torch.zeros_like(x1), torch.zeros_like(x2)
        mod.nnd_backward_cuda(x1, x2, g1, g2, gd1, gd2, i1, i2)  # __init__.py:50
        return d1, d2, i1, i2, g1, g2

    steps, warmup = args.steps, max(args.warmup, 1)
    for _ in range(warmup):
        out = step()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()  # the reference launches on the legacy default stream, which is torch's default stream
    for _ in range(steps):
        out = step()
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / steps
    pairs = 2.0 * b * n * m
    ours = F.nnd_forward_raw(x1, x2)
    same = {k: float((a == c).float().mean()) for k, a, c in zip(("dist1", "dist2", "idx1", "idx2"), out[:4], ours)}
    line = {"impl": "reference_cuda", "metric": METRIC, "value": pairs / (ms * 1e-3) / 1e9, "unit": UNIT, "n_gpus": 1,
            "steps": steps, "warmup": warmup, "ms_per_step": ms, "higher_is_better": True, "dtype": "f32",
            "data": "synthetic", "config": {"workload": w["desc"], "cloud": w["kind"], "b": b, "n": n, "m": m},
            "fraction_equal_to_b200_path": same,
            "note": "reference nnd_cuda.cu unmodified, nvcc defaults (FMA-contracted), sm_100a recompile"}
    emit(line)


# --------------------------------------------------------------------------------------------- GPU arm
def run_b200_arm(args, w):
    import torch
    import torch.distributed as dist

    import torch_nndistance as NND
    from pointcloudregistration_b200 import _cabi, sharded, synthetic
    from pointcloudregistration_b200 import functional as F

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world != args.gpus:
        raise SystemExit(f"--gpus {args.gpus} but WORLD_SIZE={world}: launch with torch.distributed.run")
    assert torch.cuda.is_available(), "bench.py needs a GPU (there is no CPU fallback)"
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    lib = _cabi.load()  # raises if libnnd_b200.so is missing
    import ctypes

    b, n, m = w["b"], w["n"], w["m"]
    pairs = 2.0 * b * n * m
    steps, warmup = args.steps, max(args.warmup, 3)
    hbm_gbs, sm_max_mhz, peak_src = load_peaks()
    fp32_peak = SMS * LANES * sm_max_mhz * 1e6  # separately-rounded FP32 ops/s (FMA-flop peak is 2x this)

    # ---- inputs: generated on the CPU from fixed seeds, pinned, then made resident
    x1h, x2h = synthetic.make_pair(w["kind"], b, n, m, w["config"])
    gd1h, gd2h = synthetic.grad_pair(b, n, m, w["config"], mean=True)
    x1h, x2h = x1h.pin_memory(), x2h.pin_memory()
    x1, x2 = x1h.to(dev), x2h.to(dev)
    gd1, gd2 = gd1h.to(dev), gd2h.to(dev)
    flush = torch.empty(256 * 1024 * 1024, dtype=torch.uint8, device=dev)  # > 126 MB L2

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    launches = [0]

    def step_resident():
        if world == 1:
            d1, d2, i1, i2 = F.nnd_forward_raw(x1, x2)
            launches[0] += lib.nnd_last_launch_count()
            g1, g2 = F.nnd_backward_raw(x1, x2, gd1, gd2, i1, i2)
            launches[0] += lib.nnd_last_launch_count()
        else:
            # rank 0's clouds are broadcast (NCCL over NVLink), rows sharded, results gathered / all-reduced
            before = sharded.CudaCompute.launches  # this rank's kernels (prepare, seed, panel scans, finalizes, grad terms)
            d1, d2, i1, i2, info = sharded.sharded_forward(x1, x2, src=0, gather=True)
            g1, g2 = sharded.sharded_backward(x1, x2, gd1, gd2, i1, i2, info, gathered=True)
            launches[0] += sharded.CudaCompute.launches - before
        return d1, d2, g1, g2

    for _ in range(warmup):
        step_resident()
    barrier()

    # ---- timed region 1: device-resident, per-step CUDA events (L2 flushed between steps, untimed)
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(steps)]
    lib.nnd_profile_enable(1)
    launches[0] = 0
    with ClockSampler(local) as clocks:
        barrier()
        for s in range(steps):
            flush.fill_(s & 0xff)
            ev[s][0].record()
            step_resident()
            ev[s][1].record()
        barrier()
    step_ms = [a.elapsed_time(b_) for a, b_ in ev]
    kernel_ms = (ctypes.c_float * 3)()
    calls = ctypes.c_int(0)
    _cabi.check(lib.nnd_profile_collect(kernel_ms, ctypes.byref(calls)), "nnd_profile_collect")
    lib.nnd_profile_enable(0)
    total_ms = torch.tensor([sum(step_ms)], dtype=torch.float64, device=dev)
    scan_ms = torch.tensor([kernel_ms[1]], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(total_ms, op=dist.ReduceOp.MAX)
        dist.all_reduce(scan_ms, op=dist.ReduceOp.MAX)
    total_ms, scan_ms_total = float(total_ms.item()), float(scan_ms.item())
    value = pairs * steps / (total_ms * 1e-3) / 1e9
    gpu_launches = launches[0]
    if world > 1:  # whole job: every rank's kernels
        lt = torch.tensor([launches[0]], dtype=torch.int64, device=dev)
        dist.all_reduce(lt, op=dist.ReduceOp.SUM)
        gpu_launches = int(lt.item())

    # backward alone (device events), for the secondary HBM roofline
    bwd_ms = None
    if world == 1:
        _, _, i1, i2 = F.nnd_forward_raw(x1, x2)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        tot = 0.0
        for s in range(max(steps, 5)):
            flush.fill_(s & 0xff)
            e0.record()
            F.nnd_backward_raw(x1, x2, gd1, gd2, i1, i2)
            e1.record()
            torch.cuda.synchronize()
            tot += e0.elapsed_time(e1)
        bwd_ms = tot / max(steps, 5)

    # ---- timed region 2: end to end through the public API, host buffers in, host buffers out
    g1h = torch.empty(b, n, 3, dtype=torch.float32).pin_memory()
    g2h = torch.empty(b, m, 3, dtype=torch.float32).pin_memory()
    h2d = (x1h.numel() + x2h.numel()) * 4
    d2h = (g1h.numel() + g2h.numel()) * 4 + 4

    def step_e2e():
        if world == 1 or rank == 0:
            a = x1h.to(dev, non_blocking=True).requires_grad_(True)
            c = x2h.to(dev, non_blocking=True).requires_grad_(True)
        else:
            a = torch.empty(b, n, 3, device=dev).requires_grad_(True)
            c = torch.empty(b, m, 3, device=dev).requires_grad_(True)
        if world == 1:
            d1, d2 = NND.nnd(a, c)
        else:
            with torch.no_grad():
                dist.broadcast(a, src=0)
                dist.broadcast(c, src=0)
            d1, d2 = sharded.nnd_sharded(a, c)
        loss = d1.mean() + d2.mean()  # chamfer as the callers form it (dip/train.py:84)
        loss.backward()
        if world == 1 or rank == 0:
            g1h.copy_(a.grad, non_blocking=True)
            g2h.copy_(c.grad, non_blocking=True)
            return loss.item()  # device -> host read of the step's result (synchronises)
        return None

    for _ in range(2):
        step_e2e()
    barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(steps):
        last_loss = step_e2e()
    e1.record()
    barrier()
    e2e_ms = torch.tensor([e0.elapsed_time(e1)], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(e2e_ms, op=dist.ReduceOp.MAX)
    e2e_value = pairs * steps / (float(e2e_ms.item()) * 1e-3) / 1e9

    # ---- CPU baseline beside it (rank 0, N=1 only)
    cpu_base = None
    if world == 1 and not args.no_cpu_baseline:
        kind = "reference" if (ROOT / "oracle" / "_ref" / "nnd_ref_cpu.so").exists() else "port"
        cpu_base, _ = cpu_arm(w, 1, 0, args.cpu_seconds, kind)

    if rank == 0:
        # One launch of the symmetric scan kernel evaluates every (xyz1[j], xyz2[k]) distance of its row range
        # ONCE and serves both directions from it (row minima -> dist1/idx1, column minima -> dist2/idx2), so a
        # launch EXECUTES b*n*m/world distance evaluations = half of the metric's `pairs` (2*b*n*m counts each
        # distance once per direction).  roofline.achieved/frac count the evaluations actually executed
        # (8 FP32 ops each) -- the honest pipe utilisation; `metric_equivalent_*` credits the launch with the
        # 2 pairs per evaluation it delivers (what the two-pass reference formulation would have had to execute).
        scan_calls_per_step = 1
        scan_ms_avg = scan_ms_total / (steps * scan_calls_per_step)  # average launch duration
        symmetric = world > 1 or bool(lib.nnd_forward_is_symmetric(n, m))
        # distance evaluations one launch on one GPU executes (small clouds take the two-pass kernel: every pair)
        scan_evals = (pairs / 2 if symmetric else pairs) / world
        achieved = scan_evals * FP32_OPS_PER_PAIR / (scan_ms_avg * 1e-3) / 1e12
        if not symmetric:
            scan_kernel_name = "nnd_scan_kernel"
        else:
            from pointcloudregistration_b200.sharded import shard_range
            lo, hi = shard_range(n, world, 0, align=lib.nnd_row_group()) if world > 1 else (0, n)
            kind = lib.nnd_rows_scan_kind(b, n, m, lo, hi)
            scan_kernel_name = {1: "nnd_scan_sym_kernel",
                                2: "nnd_scan_symc_kernel (one launch per panel of columns on rotating streams; the "
                                   "phase seed + all panel launches is timed as one)",
                                3: "nnd_scan_symc_kernel, roles swapped (row shard streamed as columns; phase timed "
                                   "as one)"}.get(kind, "nnd_scan_sym_kernel")
        peak = fp32_peak / 1e12
        traffic = traffic_note = None
        tfile = ROOT / "profiles" / "scan_traffic.json"
        if tfile.exists():
            tj = json.loads(tfile.read_text())
            traffic = tj.get(f"{w['n']}x{w['m']}x{w['b']}")
            traffic_note = tj.get("_note")
        cs = clocks.summary()
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": steps, "warmup": warmup,
            "ms_per_step": total_ms / steps, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
            "dtype": "f32", "data": "synthetic",
            "config": {"workload": w["desc"], "cloud": w["kind"], "b": b, "n": n, "m": m,
                       "pairs_per_step": pairs, "l2": "256 MB buffer rewritten between timed steps (L2 flush)",
                       "parallelism": "xyz1 rows sharded over ranks (symmetric scan: each rank evaluates its rows x all "
                                      "columns once); NCCL broadcast of both clouds, int64 MIN all_reduce of the "
                                      "column keys, all_gather of dist1/idx1, SUM all_reduce of gradients"
                                      if world > 1 else "single GPU"},
            "roofline": {"bound": "fp32", "kernel": scan_kernel_name,
                         "achieved": achieved, "peak": peak,
                         "unit": "TFLOP/s", "frac": achieved / peak, "traffic": traffic, "traffic_note": traffic_note,
                         "peak_source": f"148 SM x 128 FP32 lanes x sm_max_mhz {sm_max_mhz:.0f} MHz, {peak_src}; FMA is "
                                        "forbidden by bit-exactness so 1 op = 1 lane-slot (the FMA-flop peak is 2x)",
                         "frac_of_fma_flop_peak": achieved / (2 * peak),
                         "flop_per_launch": scan_evals * FP32_OPS_PER_PAIR, "kernel_ms": scan_ms_avg,
                         "evaluations_per_launch": scan_evals, "metric_pairs_per_launch": pairs / world,
                         "metric_equivalent_achieved": achieved * (pairs / world) / scan_evals,
                         "metric_equivalent_frac": achieved * (pairs / world) / scan_evals / peak,
                         "metric_equivalent_frac_of_fma_flop_peak": achieved * (pairs / world) / scan_evals / (2 * peak),
                         "kernel_share_of_step": scan_ms_total / world / total_ms if world == 1 else None},
            "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                    "ms_per_step": float(e2e_ms.item()) / steps, "loss": last_loss},
            "gpu_launches": gpu_launches,
            "clocks": cs,
        }
        if cs["sm_mhz"]:
            line["roofline"]["frac_at_achieved_clock"] = achieved / (SMS * LANES * cs["sm_mhz"] * 1e6 / 1e12)
        if bwd_ms is not None:
            bytes_bwd = BWD_BYTES_PER_ROW * b * (n + m)
            line["roofline_bwd"] = {"bound": "hbm", "kernel": "nnd_grad_kernel", "achieved": bytes_bwd / (bwd_ms * 1e-3) / 1e9,
                                    "peak": hbm_gbs, "unit": "GB/s", "frac": bytes_bwd / (bwd_ms * 1e-3) / 1e9 / hbm_gbs,
                                    "ms": bwd_ms, "bytes": bytes_bwd, "note": "2 memsets + 1 kernel, timed together"}
        if cpu_base is not None:
            line["cpu_baseline"] = cpu_base
        emit(line)
    if world > 1:
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", choices=["b200", "reference", "reference_cuda"], default="b200")
    ap.add_argument("--workload", choices=sorted(WORKLOADS), default="c4")
    ap.add_argument("--kind", default=None, help="cloud generator override (uniform, clustered, lattice, ...)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--cpu-seconds", type=float, default=12.0, help="CPU work in the cpu_baseline sample")
    ap.add_argument("--cpu-cores", type=int, default=0, help="cap the host cores the CPU arm uses (0 = all)")
    ap.add_argument("--cpu-seconds-per-step", type=float, default=3.0, help="--impl reference: CPU work per step")
    args = ap.parse_args()
    claim_stdout()
    w = dict(WORKLOADS[args.workload])
    cpu_arm.max_cores = args.cpu_cores or None
    if args.kind:
        w["kind"] = args.kind
    if args.impl == "reference":
        run_reference_arm(args, w)
    elif args.impl == "reference_cuda":
        run_reference_cuda_arm(args, w)
    else:
        run_b200_arm(args, w)


if __name__ == "__main__":
    main()
