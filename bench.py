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
             the FP32 issue peak 148 SM x 128 lanes x sm_max_mhz; roofline.backward = the gradient kernel vs HBM
  parity     computed OUTSIDE the timed regions from the outputs of the timed path: SHA-256 of the four forward
             tensors and of the deterministic backward's gradients against the digests the compiled reference
             produced for the same seeded inputs (tests/golden/digests_full.json), the atomic backward's error
             against the deterministic one, and -- at N > 1 -- whether every rank holds identical results
  other_configs  the remaining BASELINE.json configs (C1, the real DIP shape, C2, C3 incl. its graph-replayed loop,
             C5, C4-clustered), a few steps each, same protocol, each with its own parity digest, and
             gpu_baseline: the reference's OWN CUDA kernels (nnd_cuda.cu recompiled for sm_100a, run in a
             separate process) timed on the same inputs on the same GPU
  cpu_baseline  the reference's own my_lib.cpp (oracle/_ref, unmodified) on all host cores, on a
             bounded query sample of the same workload, same run (N=1, rank 0 only)

The oracle / reference code is executed ONLY in the cpu_baseline / gpu_baseline legs and in --impl reference*.
"""
from __future__ import annotations

import argparse
import hashlib
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

WORKLOADS = {  # BASELINE.json configs (+ the shape DIP really trains on, dip/train.py:84)
    "c1": dict(b=1, n=2048, m=2048, config=1, kind="uniform", desc="C1 B=1 N=M=2048 (DIP patch size)"),
    "dip": dict(b=256, n=256, m=256, config=6, kind="uniform", desc="DIP real shape B=256 N=M=256 (dip/train.py:84)"),
    "c2": dict(b=32, n=16384, m=16384, config=2, kind="uniform", desc="C2 DIP training-shape batch B=32 N=M=16384"),
    "c3": dict(b=1, n=100_000, m=100_000, config=3, kind="uniform", desc="C3 NDP/C2P-Net pair N=M=100k"),
    "c4": dict(b=1, n=1_000_000, m=1_000_000, config=4, kind="uniform", desc="C4 large-cloud chamfer N=M=1M"),
    "c5": dict(b=1, n=262_144, m=262_144, config=5, kind="lattice", desc="C5 tie stress N=M=262144"),
}


def digest_key(workload: str, kind: str) -> str:
    """tests/golden/digests_full.json key of a (workload, cloud kind)."""
    return f"{workload}_{kind}"


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


def load_digests():
    p = ROOT / "tests" / "golden" / "digests_full.json"
    return json.loads(p.read_text()) if p.exists() else {}


def config_of(w):
    """The workload identity, key-identical in every arm (the driver compares it between arms)."""
    return {"workload": w["desc"], "cloud": w["kind"], "b": w["b"], "n": w["n"], "m": w["m"]}


def sha_t(t) -> str:
    """SHA-256 of a tensor's bytes (device tensors are copied to the host first)."""
    return hashlib.sha256(t.detach().contiguous().cpu().numpy().tobytes()).hexdigest()


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
        "config": config_of(w),
        "config_detail": {"note": "each step is a bounded query sample of the workload on the host cores; the rate is "
                                  "per pair and extrapolates linearly (cpu_baseline.sample)"},
        "cpu_baseline": base,
        "e2e": {"value": base["value"], "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    emit(line)


# --------------------------------------------------------------------------------------------- reference GPU kernel
def time_reference_cuda(w, steps, warmup):
    """The reference's OWN CUDA kernels (nnd_cuda.cu, unmodified, recompiled for sm_100a into
    oracle/_ref/nnd_ref_cuda.so) driven the way torch_nndistance/__init__.py:11-57 drives them, same metric, same
    inputs, 1 GPU.  Not a parity oracle (FMA-contracted): the record says how far its outputs are from the B200
    path's (which is bit-exact against the reference CPU code)."""
    import importlib.util

    import torch

    from pointcloudregistration_b200 import functional as F
    from pointcloudregistration_b200 import synthetic

    so = ROOT / "oracle" / "_ref" / "nnd_ref_cuda.so"
    if not so.exists():
        return {"unavailable": "oracle/_ref/nnd_ref_cuda.so not built"}
    mod = sys.modules.get("nnd_ref_cuda")
    if mod is None:
        spec = importlib.util.spec_from_file_location("nnd_ref_cuda", str(so))
        mod = importlib.util.module_from_spec(spec)
        spec.loader.exec_module(mod)
        sys.modules["nnd_ref_cuda"] = mod
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

    out = None
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
    return {"value": pairs / (ms * 1e-3) / 1e9, "unit": UNIT, "ms_per_step": ms, "steps": steps, "warmup": warmup,
            "fraction_equal_to_b200_path": same}


def run_reference_cuda_arm(args, w):
    """--impl reference_cuda: one line for --workload, or (with --workloads a,b,...) one line holding a record per
    workload -- the form the default arm's gpu_baseline leg runs in a separate process."""
    if int(os.environ.get("RANK", "0")) != 0:
        return
    note = "reference nnd_cuda.cu unmodified, nvcc defaults (FMA-contracted), sm_100a recompile"
    if args.workloads:
        recs = {}
        for name in args.workloads.split(","):
            key, _, kind = name.partition(":")
            ww = dict(WORKLOADS[key])
            if kind:
                ww["kind"] = kind
            big = ww["b"] * ww["n"] * ww["m"] >= 5e11  # ~6 s per step at 1M x 1M: one warm-up, one timed step
            tiny = ww["b"] * ww["n"] * ww["m"] <= 1e9
            recs[name] = time_reference_cuda(ww, 1 if big else 50 if tiny else 3, 1 if big else 5 if tiny else 2)
        emit({"impl": "reference_cuda", "note": note, "records": recs})
        return
    rec = time_reference_cuda(w, args.steps, max(args.warmup, 1))
    if "unavailable" in rec:
        emit({"impl": "reference_cuda", **rec})
        return
    emit({"impl": "reference_cuda", "metric": METRIC, "value": rec["value"], "unit": UNIT, "n_gpus": 1,
          "steps": rec["steps"], "warmup": rec["warmup"], "ms_per_step": rec["ms_per_step"], "higher_is_better": True,
          "dtype": "f32", "data": "synthetic", "config": config_of(w),
          "fraction_equal_to_b200_path": rec["fraction_equal_to_b200_path"], "note": note})


# --------------------------------------------------------------------------------------------- GPU arm
SCAN_NAMES = {0: "nnd_small_kernel (single launch: whole target cloud in shared memory, index resolved in-kernel)",
              1: "nnd_scan_kernel (two-pass, TMA ring)",
              "sym": "nnd_scan_sym_kernel (persistent, TMA ring; every distance evaluated once)",
              "panel": "nnd_scan_symc_kernel (one launch per panel of columns on rotating streams; the phase seed + all "
                       "panel launches is timed as one)",
              "swapped": "nnd_scan_symc_kernel, roles swapped (row shard streamed as columns; phase timed as one)"}


class _Null:
    def __enter__(self):
        return self

    def __exit__(self, *exc):
        return False


class Runner:
    """One workload resident on this rank's GPU: the timed step, the roofline of its scan, and its parity record."""

    def __init__(self, w, world, rank, dev, lib, fp32_peak, digests):
        import torch

        from pointcloudregistration_b200 import synthetic

        self.torch, self.w, self.world, self.rank, self.dev, self.lib = torch, w, world, rank, dev, lib
        self.fp32_peak, self.digests = fp32_peak, digests
        b, n, m = w["b"], w["n"], w["m"]
        self.b, self.n, self.m = b, n, m
        self.pairs = 2.0 * b * n * m
        self.x1h, self.x2h = synthetic.make_pair(w["kind"], b, n, m, w["config"])
        gd1h, gd2h = synthetic.grad_pair(b, n, m, w["config"], mean=True)
        self.x1h, self.x2h = self.x1h.pin_memory(), self.x2h.pin_memory()
        self.x1, self.x2 = self.x1h.to(dev), self.x2h.to(dev)
        self.gd1, self.gd2 = gd1h.to(dev), gd2h.to(dev)
        self.launches = 0
        self.last = None

    def step(self):
        """forward + backward with the inputs resident; keeps the outputs in self.last."""
        from pointcloudregistration_b200 import functional as F
        from pointcloudregistration_b200 import sharded

        lib = self.lib
        if self.world == 1:
            d1, d2, i1, i2 = F.nnd_forward_raw(self.x1, self.x2)
            self.launches += lib.nnd_last_launch_count()
            g1, g2 = F.nnd_backward_raw(self.x1, self.x2, self.gd1, self.gd2, i1, i2)
            self.launches += lib.nnd_last_launch_count()
        else:
            # rank 0's clouds are broadcast (NCCL over NVLink), rows sharded, results gathered / all-reduced
            before = sharded.CudaCompute.launches
            d1, d2, i1, i2, info = sharded.sharded_forward(self.x1, self.x2, src=0, gather=True)
            g1, g2 = sharded.sharded_backward(self.x1, self.x2, self.gd1, self.gd2, i1, i2, info, gathered=True)
            self.launches += sharded.CudaCompute.launches - before
        self.last = (d1, d2, i1, i2, g1, g2)

    def timed(self, steps, warmup, flush, barrier, sampler=None):
        """(step_ms list, scan ms per profiled call, profiled calls): per-step CUDA events on the launching stream, L2
        flushed between steps (untimed), the library's own events around the scan inside the same steps."""
        import ctypes

        from pointcloudregistration_b200 import _cabi

        torch, lib = self.torch, self.lib
        for _ in range(warmup):
            self.step()
        barrier()
        ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(steps)]
        scan_total, calls_total = 0.0, 0
        self.launches = 0
        with (sampler if sampler is not None else _Null()):
            barrier()
            done = 0
            while done < steps:  # the library remembers 64 profiled calls: collect in chunks, never drop one
                chunk = min(steps - done, 48)
                lib.nnd_profile_enable(1)
                for s in range(done, done + chunk):
                    flush.fill_(s & 0xff)
                    ev[s][0].record()
                    self.step()
                    ev[s][1].record()
                kernel_ms = (ctypes.c_float * 3)()
                calls = ctypes.c_int(0)
                _cabi.check(lib.nnd_profile_collect(kernel_ms, ctypes.byref(calls)), "nnd_profile_collect")
                lib.nnd_profile_enable(0)
                assert calls.value == chunk, f"profiler covered {calls.value} scans for {chunk} steps"
                scan_total += kernel_ms[1]
                calls_total += calls.value
                done += chunk
            barrier()
        step_ms = [a.elapsed_time(b_) for a, b_ in ev]
        return step_ms, scan_total / max(calls_total, 1), calls_total

    def scan_kernel(self):
        """(name, distance evaluations one rank's scan executes per step)"""
        from pointcloudregistration_b200.sharded import uniform_shard_range

        lib, b, n, m, world = self.lib, self.b, self.n, self.m, self.world
        if world == 1:
            kind = lib.nnd_forward_kind(b, n, m)
            if kind in (0, 1):
                return SCAN_NAMES[kind], self.pairs
            lo, hi = 0, n
        else:
            lo, hi = uniform_shard_range(n, world, 0, align=lib.nnd_row_group())
        rk = lib.nnd_rows_scan_kind(b, n, m, lo, hi)
        return SCAN_NAMES[{1: "sym", 2: "panel", 3: "swapped"}.get(rk, "sym")], self.pairs / 2 / world

    def roofline(self, scan_ms):
        name, evals = self.scan_kernel()
        achieved = evals * FP32_OPS_PER_PAIR / (scan_ms * 1e-3) / 1e12
        peak = self.fp32_peak / 1e12
        return {"bound": "fp32", "kernel": name, "achieved": achieved, "peak": peak, "unit": "TFLOP/s",
                "frac": achieved / peak, "kernel_ms": scan_ms, "evaluations_per_launch": evals,
                "flop_per_launch": evals * FP32_OPS_PER_PAIR, "metric_pairs_per_launch": self.pairs / self.world,
                "metric_equivalent_frac": achieved * (self.pairs / self.world) / evals / peak}

    def parity(self, name):
        """Outside every timed region.  Compares the LAST timed step's outputs (what the timed path produced) with the
        compiled reference's digests, runs the deterministic backward once for the gradient digests, and measures the
        atomic backward (the one that was timed) against it."""
        import torch.distributed as dist

        from pointcloudregistration_b200 import functional as F
        from pointcloudregistration_b200 import sharded

        torch = self.torch
        d1, d2, i1, i2, g1, g2 = self.last
        rec = {"digest_case": name,
               "source": "tests/golden/digests_full.json (compiled reference my_lib.cpp, whole tensors)"}
        got = {k: sha_t(v) for k, v in zip(("dist1", "dist2", "idx1", "idx2"), (d1, d2, i1, i2))}
        meta = self.digests.get(name)
        if meta is None:
            rec["forward"] = "no committed digest for this case"
        else:
            bad = [k for k in got if got[k] != meta["sha256"][k]]
            rec["forward"] = "bit-exact" if not bad else "MISMATCH: " + ",".join(bad)
        if self.world == 1:
            e1, e2 = F.nnd_backward_raw(self.x1, self.x2, self.gd1, self.gd2, i1, i2, exact=True)
        else:
            info = sharded.ShardInfo(self.rank, self.world, 0, 0, 0, 0)
            e1, e2 = sharded.sharded_backward(self.x1, self.x2, self.gd1, self.gd2, i1, i2, info, gathered=True,
                                              exact=True)
        gotg = {"grad1_mean": sha_t(e1), "grad2_mean": sha_t(e2)}
        if meta is not None and "grad1_mean" in meta["sha256"]:
            badg = [k for k in gotg if gotg[k] != meta["sha256"][k]]
            rec["backward_deterministic"] = "bit-exact" if not badg else "MISMATCH: " + ",".join(badg)
        scale = max(float(e1.abs().max()), float(e2.abs().max()), 1e-30)
        rec["grad_max_rel_err"] = max(float((g1 - e1).abs().max()), float((g2 - e2).abs().max())) / scale
        rec["grad_tolerance"] = 1e-5
        if self.world > 1:
            mine = bytes.fromhex("".join(got[k] for k in ("dist1", "dist2", "idx1", "idx2"))
                                 + gotg["grad1_mean"] + gotg["grad2_mean"])
            t = torch.tensor(list(mine), dtype=torch.uint8, device=self.dev)
            allt = [torch.empty_like(t) for _ in range(self.world)]
            dist.all_gather(allt, t)
            rec["ranks_identical"] = all(bool(torch.equal(a, allt[0])) for a in allt)
            rec["ranks"] = self.world
        return rec


def graph_step_ms(r, reps=20):
    """Device time of one forward+backward free of host launch gaps: a CUDA graph of `reps` steps over static buffers
    through the C ABI (latency-bound shapes only: the eager number is bounded by Python, not by the GPU)."""
    torch, lib = r.torch, r.lib
    b, n, m = r.b, r.n, r.m
    dev = r.dev
    d1, d2 = torch.empty(b, n, device=dev), torch.empty(b, m, device=dev)
    i1 = torch.empty(b, n, dtype=torch.int32, device=dev)
    i2 = torch.empty(b, m, dtype=torch.int32, device=dev)
    g1, g2 = torch.empty_like(r.x1), torch.empty_like(r.x2)
    ws = torch.empty(max(lib.nnd_workspace_bytes(b, n, m), 256), dtype=torch.uint8, device=dev)

    def body():
        st = torch.cuda.current_stream().cuda_stream
        rc = lib.nnd_forward_cuda_f32(r.x1.data_ptr(), r.x2.data_ptr(), b, n, m, d1.data_ptr(), d2.data_ptr(),
                                      i1.data_ptr(), i2.data_ptr(), ws.data_ptr(), ws.numel(), st)
        rc = rc or lib.nnd_backward_cuda_f32(r.x1.data_ptr(), r.x2.data_ptr(), r.gd1.data_ptr(), r.gd2.data_ptr(),
                                             i1.data_ptr(), i2.data_ptr(), b, n, m, g1.data_ptr(), g2.data_ptr(), st)
        assert rc == 0, rc

    s = torch.cuda.Stream()
    with torch.cuda.stream(s):
        body()
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        for _ in range(reps):
            body()
    g.replay()
    torch.cuda.synchronize()
    times = []
    for _ in range(5):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        g.replay()
        e1.record()
        torch.cuda.synchronize()
        times.append(e0.elapsed_time(e1) / reps)
    return statistics.median(times)


def e2e_region(r, steps, warmup, barrier):
    """The step through the public API (torch_nndistance.nnd + autograd) from PINNED HOST buffers: H2D of both clouds,
    D2H of both gradients and the loss inside the timed region, every step.  (ms per step, last loss, h2d, d2h)."""
    import torch.distributed as dist

    import torch_nndistance as NND
    from pointcloudregistration_b200 import sharded

    torch, dev, world, rank = r.torch, r.dev, r.world, r.rank
    b, n, m = r.b, r.n, r.m
    g1h = torch.empty(b, n, 3, dtype=torch.float32).pin_memory()
    g2h = torch.empty(b, m, 3, dtype=torch.float32).pin_memory()
    h2d = (r.x1h.numel() + r.x2h.numel()) * 4
    d2h = (g1h.numel() + g2h.numel()) * 4 + 4

    def step_e2e():
        # every rank holds the step's inputs in its own pinned host memory (one process per GPU, SPMD) and uploads them
        # itself, in parallel over its own PCIe link: no rank-0 upload followed by an NVLink broadcast
        a = r.x1h.to(dev, non_blocking=True).requires_grad_(True)
        c = r.x2h.to(dev, non_blocking=True).requires_grad_(True)
        if world == 1:
            d1, d2 = NND.nnd(a, c)
        else:
            d1, d2 = sharded.nnd_sharded(a, c)
        loss = d1.mean() + d2.mean()  # chamfer as the callers form it (dip/train.py:84)
        loss.backward()
        if world == 1 or rank == 0:
            g1h.copy_(a.grad, non_blocking=True)
            g2h.copy_(c.grad, non_blocking=True)
            return loss.item()  # device -> host read of the step's result (synchronises)
        return None

    last = None
    for _ in range(warmup):
        step_e2e()
    barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(steps):
        last = step_e2e()
    e1.record()
    barrier()
    ms = torch.tensor([e0.elapsed_time(e1)], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    return float(ms.item()) / steps, last, h2d, d2h


def e2e_graph_region(r, steps, warmup):
    """End to end through the repo's OTHER public entry point for repeated chamfer on fixed-size clouds,
    ndp_loop.ChamferLoop (forward + mean/mean chamfer epilogue + backward replayed as one CUDA graph): every step
    copies both clouds from PINNED HOST buffers into the loop's static inputs, replays, and copies both gradients and
    the loss back to the host.  Same work and same bytes as e2e_region, without autograd's per-step host cost."""
    from pointcloudregistration_b200 import ndp_loop

    torch, dev = r.torch, r.dev
    b, n, m = r.b, r.n, r.m
    loop = ndp_loop.ChamferLoop(b, n, m, dev, trunc=float("inf"))
    g1h = torch.empty(b, n, 3, dtype=torch.float32).pin_memory()
    g2h = torch.empty(b, m, 3, dtype=torch.float32).pin_memory()
    lossh = torch.empty((), dtype=torch.float32).pin_memory()

    def step():
        loop.y.copy_(r.x2h, non_blocking=True)
        loop.step(r.x1h)  # copies x (pinned host -> static device buffer) and replays the graph
        g1h.copy_(loop.grad_x, non_blocking=True)
        g2h.copy_(loop.grad_y, non_blocking=True)
        lossh.copy_(loop.loss, non_blocking=True)
        torch.cuda.current_stream().synchronize()
        return float(lossh)

    last = None
    for _ in range(warmup):
        step()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(steps):
        last = step()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / steps, last


def other_config(key, kind, world, rank, dev, lib, fp32_peak, digests, flush, barrier, want_e2e, want_graph):
    """One of the non-headline configs: a few timed steps with the headline's protocol, its roofline fraction and its
    own parity record."""
    import torch
    import torch.distributed as dist

    w = dict(WORKLOADS[key])
    if kind:
        w["kind"] = kind
    r = Runner(w, world, rank, dev, lib, fp32_peak, digests)
    tiny = r.pairs <= 1e9
    steps = 40 if tiny else 3
    step_ms, scan_ms, _ = r.timed(steps, 5 if tiny else 3, flush, barrier)
    total = torch.tensor([sum(step_ms)], dtype=torch.float64, device=dev)
    scan = torch.tensor([scan_ms], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(total, op=dist.ReduceOp.MAX)
        dist.all_reduce(scan, op=dist.ReduceOp.MAX)
    ms = float(total.item()) / steps
    rec = {"config": config_of(w), "steps": steps, "ms_per_step": ms, "value": r.pairs / (ms * 1e-3) / 1e9, "unit": UNIT,
           "gpu_launches_per_step": r.launches / steps,
           "roofline": r.roofline(float(scan.item())),
           "parity": r.parity(digest_key(key, w["kind"]))}
    if want_graph and world == 1:
        gms = graph_step_ms(r)
        rec["graph_replay"] = {"ms_per_step": gms, "value": r.pairs / (gms * 1e-3) / 1e9, "unit": UNIT,
                               "note": "CUDA graph of 20 forward+backward pairs through the C ABI: device time without "
                                       "host launch gaps (ms_per_step above is the eager Python path, host-bound here)"}
    if want_e2e:
        ems, loss, h2d, d2h = e2e_region(r, 40 if tiny else 3, 5 if tiny else 2, barrier)
        rec["e2e"] = {"value": r.pairs / (ems * 1e-3) / 1e9, "unit": UNIT, "ms_per_step": ems, "loss": loss,
                      "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h, "api": "torch_nndistance.nnd + autograd"}
        if world == 1:
            gms, gloss = e2e_graph_region(r, 40 if tiny else 3, 5 if tiny else 2)
            rec["e2e_graph"] = {"value": r.pairs / (gms * 1e-3) / 1e9, "unit": UNIT, "ms_per_step": gms, "loss": gloss,
                                "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                                "api": "ndp_loop.ChamferLoop (one CUDA-graph replay per step; host buffers in and out)"}
    return rec, r


def pruned_leg(dev, lib, fp32_peak, digests, flush, brute_ms):
    """The opt-in exact pruned search (nnd_forward_pruned_cuda_f32, csrc/nnd_pruned.cuh) beside the brute-force step:
    forward (pruned) + backward (the same kernel as everywhere) with the headline's protocol -- inputs resident, L2
    flushed between steps, CUDA events -- and the forward outputs held to the SAME full-size reference digests.
    `value` is the metric the brute-force step is quoted in (2*B*N*M nominal pairs / time): not a pipe utilisation, the
    search evaluates only the pairs that can matter.  brute_ms: ms_per_step of the brute-force step of each case."""
    import torch

    from pointcloudregistration_b200 import functional as F

    out = {}
    for key, kind in (("c2", None), ("c3", None), ("c5", None), ("c4", "uniform"), ("c4", "clustered")):
        w = dict(WORKLOADS[key])
        if kind:
            w["kind"] = kind
        name = key if not kind or kind == WORKLOADS[key]["kind"] else f"{key}:{kind}"
        r = Runner(w, 1, 0, dev, lib, fp32_peak, digests)
        launches = 0

        def step():
            nonlocal launches
            d1, d2, i1, i2 = F.nnd_forward_pruned_raw(r.x1, r.x2)
            launches = lib.nnd_last_launch_count()
            g1, g2 = F.nnd_backward_raw(r.x1, r.x2, r.gd1, r.gd2, i1, i2)
            return d1, d2, i1, i2

        for _ in range(3):
            last = step()
        torch.cuda.synchronize()
        steps = 5
        ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(steps)]
        for i in range(steps):
            flush.fill_(i & 0xff)
            ev[i][0].record()
            last = step()
            ev[i][1].record()
        torch.cuda.synchronize()
        ms = sum(a.elapsed_time(b_) for a, b_ in ev) / steps
        meta = digests.get(digest_key(key, w["kind"]))
        if meta is None:
            fwd = "no committed digest for this case"
        else:
            bad = [k for k, v in zip(("dist1", "dist2", "idx1", "idx2"), last) if sha_t(v) != meta["sha256"][k]]
            fwd = "bit-exact" if not bad else "MISMATCH: " + ",".join(bad)
        rec = {"config": config_of(w), "steps": steps, "ms_per_step": ms, "value": r.pairs / (ms * 1e-3) / 1e9,
               "unit": UNIT, "forward": fwd, "own_kernels_per_forward": launches}
        if brute_ms.get(name):
            rec["brute_force_ms_per_step"] = brute_ms[name]
            rec["speedup_vs_brute_force"] = brute_ms[name] / ms
        if key == "c3":  # the same 100-iteration graph loop as other_configs.c3.loop, pruned forward inside the graph
            rec["loop"] = c3_loop(r, search="pruned")
        out[name] = rec
        del r, last
        torch.cuda.empty_cache()
    return {"api": "nnd_forward_pruned_cuda_f32 (opt-in; functional.set_search_mode('pruned')) + nnd_backward_cuda_f32",
            "note": "exact search with spatial pruning: same bits as the brute-force scan (digest source: "
                    "tests/golden/digests_full.json, compiled reference), only the pairs that can matter are evaluated; "
                    "NOT the headline path -- the metric and the roofline above are defined on the brute-force scan",
            "records": out}


def c3_loop(r, iters=100, search="brute"):
    """BASELINE.json configs[2]: the repeated-chamfer loop on the C3 pair as ndp_loop.ChamferLoop runs it (forward,
    fused mean/mean epilogue, backward captured in one CUDA graph; cloud 1 follows its gradient; loss read every 10th
    iteration as registration.py:250-256 needs).  ms per iteration from CUDA events around the whole loop."""
    from pointcloudregistration_b200 import ndp_loop

    torch = r.torch
    loop = ndp_loop.ChamferLoop(r.b, r.n, r.m, r.dev, search=search)
    loop.set_target(r.x2)
    x = r.x1.clone()
    for _ in range(10):  # untimed: graph capture and upload, clocks
        _, g = loop.step(x)
        x = x - 0.01 * r.n * g
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    first = last = None
    for it in range(iters):
        _, g = loop.step(x)
        if it % 10 == 0:
            last = loop.loss_value()
            first = last if first is None else first
        x = x - 0.01 * r.n * g
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / iters
    return {"iterations": iters, "ms_per_iteration": ms, "value": r.pairs / (ms * 1e-3) / 1e9, "unit": UNIT,
            "loss_first": first, "loss_last": last,
            "search": search,
            "note": "ndp_loop.ChamferLoop: forward + fused truncated-chamfer epilogue + backward replayed as one CUDA "
                    "graph, x <- x - 0.01 * d(sum)/dx between iterations, loss read every 10th iteration"}


def gpu_baseline_leg(names):
    """The reference's own CUDA kernels on the same inputs, same GPU, in a SEPARATE process (this process never loads
    anything under oracle/)."""
    cmd = [sys.executable, str(ROOT / "bench.py"), "--impl", "reference_cuda", "--workloads", ",".join(names)]
    try:
        p = subprocess.run(cmd, capture_output=True, text=True, timeout=600)
        if p.returncode != 0:
            return {"unavailable": f"reference_cuda leg failed: {p.stderr[-300:]}"}
        return json.loads(p.stdout.strip().splitlines()[-1])
    except Exception as exc:  # noqa: BLE001
        return {"unavailable": f"{type(exc).__name__}: {str(exc)[:200]}"}


def head_name_of(workload, w):
    return workload if w["kind"] == WORKLOADS[workload]["kind"] else f"{workload}:{w['kind']}"


def run_b200_arm(args, w):
    import torch
    import torch.distributed as dist

    from pointcloudregistration_b200 import _cabi

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

    steps, warmup = args.steps, max(args.warmup, 3)
    hbm_gbs, sm_max_mhz, peak_src = load_peaks()
    fp32_peak = SMS * LANES * sm_max_mhz * 1e6  # separately-rounded FP32 ops/s (FMA-flop peak is 2x this)
    digests = load_digests()
    flush = torch.empty(256 * 1024 * 1024, dtype=torch.uint8, device=dev)  # > 126 MB L2

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    r = Runner(w, world, rank, dev, lib, fp32_peak, digests)
    b, n, m, pairs = r.b, r.n, r.m, r.pairs

    # ---- timed region 1: device-resident, per-step CUDA events (L2 flushed between steps, untimed)
    clocks = ClockSampler(local)
    step_ms, scan_ms_avg, calls = r.timed(steps, warmup, flush, barrier, sampler=clocks)
    total_ms = torch.tensor([sum(step_ms)], dtype=torch.float64, device=dev)
    scan_ms = torch.tensor([scan_ms_avg], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(total_ms, op=dist.ReduceOp.MAX)
        dist.all_reduce(scan_ms, op=dist.ReduceOp.MAX)
    total_ms, scan_ms_avg = float(total_ms.item()), float(scan_ms.item())
    value = pairs * steps / (total_ms * 1e-3) / 1e9
    gpu_launches = r.launches
    if world > 1:  # whole job: every rank's kernels
        lt = torch.tensor([r.launches], dtype=torch.int64, device=dev)
        dist.all_reduce(lt, op=dist.ReduceOp.SUM)
        gpu_launches = int(lt.item())

    # ---- parity of what was just timed (outside the timed region)
    parity = r.parity(digest_key(args.workload, w["kind"]))

    # backward alone (device events), for the secondary HBM roofline
    bwd_ms = None
    if world == 1:
        from pointcloudregistration_b200 import functional as F

        i1, i2 = r.last[2], r.last[3]
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        tot, reps = 0.0, max(steps, 5)
        for s in range(reps):
            flush.fill_(s & 0xff)
            e0.record()
            F.nnd_backward_raw(r.x1, r.x2, r.gd1, r.gd2, i1, i2)
            e1.record()
            torch.cuda.synchronize()
            tot += e0.elapsed_time(e1)
        bwd_ms = tot / reps

    # ---- timed region 2: end to end through the public API, host buffers in, host buffers out
    e2e_ms, last_loss, h2d, d2h = e2e_region(r, steps, 2, barrier)
    e2e_value = pairs / (e2e_ms * 1e-3) / 1e9

    # ---- the other configs (a few steps each, same protocol, own parity)
    others = {}
    if not args.no_other_configs:
        other_kind = "clustered" if w["kind"] != "clustered" else "uniform"
        plan = ([("c1", None, True, True), ("dip", None, True, True), ("c2", None, False, False),
                 ("c3", None, False, False), ("c5", None, False, False), ("c4", other_kind, False, False)]
                if world == 1 else [("c4", other_kind, False, False)])
        for key, kind, want_e2e, want_graph in plan:
            if key == args.workload and (kind or WORKLOADS[key]["kind"]) == w["kind"]:
                continue
            rec, rr = other_config(key, kind, world, rank, dev, lib, fp32_peak, digests, flush, barrier, want_e2e,
                                   want_graph)
            if key == "c3" and world == 1:
                rec["loop"] = c3_loop(rr)
            others[key if not kind else f"{key}:{kind}"] = rec
            del rr
            torch.cuda.empty_cache()

    # ---- the opt-in exact pruned search beside the brute-force steps (N=1, default workload only)
    pruned = None
    if world == 1 and not args.no_other_configs and args.workload == "c4":
        brute_ms = {k: v["ms_per_step"] for k, v in others.items()}
        brute_ms[head_name_of(args.workload, w)] = total_ms / steps
        pruned = pruned_leg(dev, lib, fp32_peak, digests, flush, brute_ms)

    # ---- baselines beside it (N=1 only): the reference's CPU code on the host cores, its CUDA kernels on this GPU
    cpu_base = gpu_base = None
    head_name = args.workload if w["kind"] == WORKLOADS[args.workload]["kind"] else f"{args.workload}:{w['kind']}"
    if world == 1 and not args.no_cpu_baseline:
        kind = "reference" if (ROOT / "oracle" / "_ref" / "nnd_ref_cpu.so").exists() else "port"
        cpu_base, _ = cpu_arm(w, 1, 0, args.cpu_seconds, kind)
    if world == 1 and not args.no_gpu_baseline:
        gpu_base = gpu_baseline_leg([head_name] + list(others))
        for k, rec in others.items():
            ref = (gpu_base.get("records") or {}).get(k)
            if ref and "ms_per_step" in ref:
                rec["reference_cuda_ms_per_step"] = ref["ms_per_step"]
                rec["speedup_vs_reference_cuda"] = ref["ms_per_step"] / rec["ms_per_step"]
                for kk in ("e2e", "e2e_graph"):  # the reference's step has no host copies in it: a conservative ratio
                    if kk in rec:
                        rec[kk]["speedup_vs_reference_cuda_resident_step"] = ref["ms_per_step"] / rec[kk]["ms_per_step"]

    if rank == 0:
        # One launch of the symmetric scan kernel evaluates every (xyz1[j], xyz2[k]) distance of its row range
        # ONCE and serves both directions from it (row minima -> dist1/idx1, column minima -> dist2/idx2), so a
        # launch EXECUTES b*n*m/world distance evaluations = half of the metric's `pairs` (2*b*n*m counts each
        # distance once per direction).  roofline.achieved/frac count the evaluations actually executed
        # (8 FP32 ops each) -- the honest pipe utilisation; `metric_equivalent_*` credits the launch with the
        # 2 pairs per evaluation it delivers (what the two-pass reference formulation would have had to execute).
        roof = r.roofline(scan_ms_avg)
        peak = roof["peak"]
        traffic = traffic_note = None
        tfile = ROOT / "profiles" / "scan_traffic.json"
        if tfile.exists():
            tj = json.loads(tfile.read_text())
            traffic = tj.get(f"{n}x{m}x{b}")
            traffic_note = tj.get("_note")
        cs = clocks.summary()
        roof.update({
            "traffic": traffic, "traffic_note": traffic_note,
            "peak_source": f"148 SM x 128 FP32 lanes x sm_max_mhz {sm_max_mhz:.0f} MHz, {peak_src}; FMA is "
                           "forbidden by bit-exactness so 1 op = 1 lane-slot (the FMA-flop peak is 2x)",
            "frac_of_fma_flop_peak": roof["achieved"] / (2 * peak),
            "metric_equivalent_achieved": roof["metric_equivalent_frac"] * peak,
            "metric_equivalent_frac_of_fma_flop_peak": roof["metric_equivalent_frac"] / 2,
            "profiled_scan_calls": calls,
            "kernel_share_of_step": scan_ms_avg * steps / total_ms})
        if cs["sm_mhz"]:
            roof["frac_at_achieved_clock"] = roof["achieved"] / (SMS * LANES * cs["sm_mhz"] * 1e6 / 1e12)
        if bwd_ms is not None:
            bytes_bwd = BWD_BYTES_PER_ROW * b * (n + m)
            roof["backward"] = {"bound": "hbm", "kernel": "nnd_grad_kernel",
                                "achieved": bytes_bwd / (bwd_ms * 1e-3) / 1e9, "peak": hbm_gbs, "unit": "GB/s",
                                "frac": bytes_bwd / (bwd_ms * 1e-3) / 1e9 / hbm_gbs, "ms": bwd_ms, "bytes": bytes_bwd,
                                "note": "zero-init + gradient kernel, timed together"}
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": steps, "warmup": warmup,
            "ms_per_step": total_ms / steps, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
            "dtype": "f32", "data": "synthetic",
            "config": config_of(w),
            "config_detail": {"pairs_per_step": pairs, "l2": "256 MB buffer rewritten between timed steps (L2 flush)",
                              "parallelism": "xyz1 rows sharded over ranks (symmetric scan: each rank evaluates its rows x "
                                             "all columns once); NCCL broadcast of both clouds; exchange over peer memory "
                                             "(NVLink P2P, torch symmetric memory): one kernel merges the column keys, "
                                             "finalizes the own slice and stores dist/idx into every rank, one kernel sums "
                                             "the gradients (NCCL int64 MIN all_reduce + all_gather + SUM all_reduce when "
                                             "symmetric memory is unavailable or NND_B200_PEER_EXCHANGE=0); "
                                             "e2e: every rank uploads the inputs from its own pinned host copy (no "
                                             "broadcast), rank 0 downloads the gradients and the loss"
                                             if world > 1 else "single GPU"},
            "roofline": roof,
            "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                    "ms_per_step": e2e_ms, "loss": last_loss},
            "parity": parity,
            "gpu_launches": gpu_launches,
            "clocks": cs,
        }
        if others:
            line["other_configs"] = others
        if pruned is not None:
            line["pruned_search"] = pruned
        if cpu_base is not None:
            line["cpu_baseline"] = cpu_base
        if gpu_base is not None:
            line["gpu_baseline"] = gpu_base
            head = (gpu_base.get("records") or {}).get(head_name)
            if head and "ms_per_step" in head:
                line["speedup_vs_reference_cuda"] = head["ms_per_step"] / (total_ms / steps)
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
    ap.add_argument("--workloads", default=None, help="--impl reference_cuda: comma list of workload[:cloud] to time")
    ap.add_argument("--kind", default=None, help="cloud generator override (uniform, clustered, lattice, ...)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-gpu-baseline", action="store_true")
    ap.add_argument("--no-other-configs", action="store_true")
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
