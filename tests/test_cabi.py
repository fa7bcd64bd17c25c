"""CPU tests: libnnd_b200.so builds for sm_100a, loads, and exports every symbol include/nnd_b200.h
declares.  No compute calls (there is no GPU here)."""
import ctypes
import re
import subprocess
from pathlib import Path

import pytest

ROOT = Path(__file__).resolve().parent.parent
HEADER = ROOT / "include" / "nnd_b200.h"


def declared_symbols():
    text = re.sub(r"/\*.*?\*/", "", HEADER.read_text(), flags=re.S)
    return sorted(set(re.findall(r"\b(nnd_[a-z0-9_]+)\s*\(", text)))


def test_header_declares_the_expected_entry_points():
    syms = declared_symbols()
    for s in ("nnd_forward_cuda_f32", "nnd_backward_cuda_f32", "nnd_query_cuda_f32", "nnd_backward_term_cuda_f32",
              "nnd_workspace_bytes", "nnd_version", "nnd_forward_rows_cuda_f32", "nnd_finalize_cols_cuda_f32"):
        assert s in syms


def test_library_loads_and_exports_every_declared_symbol(cuda_lib):
    from pointcloudregistration_b200 import _cabi

    raw = ctypes.CDLL(str(_cabi.LIB_PATH))
    for s in declared_symbols():
        assert hasattr(raw, s), f"{s} declared in include/nnd_b200.h but not exported"
    assert set(_cabi.SIGNATURES) == set(declared_symbols()), "ctypes signature table out of sync with the header"


def test_version_and_error_strings(cuda_lib):
    assert b"sm_100a" in cuda_lib.nnd_version()
    assert cuda_lib.nnd_error_string(0) == b"success"
    assert b"workspace" in cuda_lib.nnd_error_string(10002)


def test_workspace_size_formula(cuda_lib):
    # SoA copies (12 B/point, padded to 512-point tiles) + 8 B/point keys for both clouds + 4 B/point column filter,
    # + 8 B/point for the second key array of a role-swapped row range, + the automatic-hint scratch (4 B/point of hints
    # and a hash table of 2^22 slots for a million rows)
    w = cuda_lib.nnd_workspace_bytes(1, 1_000_000, 1_000_000)
    assert 1_000_000 * 52 + 4_000_000 + 4 * 2 ** 22 <= w <= 1_000_000 * 52 + 4_000_000 + 4 * 2 ** 22 + 64 * 1024
    assert cuda_lib.nnd_workspace_bytes(0, 10, 10) == 256
    # the row-range entry point keeps the column keys in a caller-owned buffer (they are all-reduced)
    assert cuda_lib.nnd_rows_workspace_bytes(1, 1_000_000, 1_000_000) + 8 * cuda_lib.nnd_colkeys_count(1, 1_000_000) <= w
    assert cuda_lib.nnd_colkeys_count(3, 513) == 3 * 1024 and cuda_lib.nnd_row_group() == 8
    assert cuda_lib.nnd_query_workspace_bytes(4, 5120, 5120) == 2 * (4 * 5120 * 12) + 4 * 5120 * 8
    # small clouds take the single-launch kernels: no workspace at all
    assert cuda_lib.nnd_small_max_points() == 4096
    assert cuda_lib.nnd_workspace_bytes(256, 256, 256) == 0 and cuda_lib.nnd_query_workspace_bytes(2, 512, 512) == 0
    assert cuda_lib.nnd_forward_kind(256, 256, 256) == 0 and cuda_lib.nnd_forward_kind(4, 4096, 4097) == 1
    assert cuda_lib.nnd_forward_kind(32, 16384, 16384) == 2 and cuda_lib.nnd_ragged_supported(3000, 2900) == 1
    # a lone pair keeps the single launch up to ~10 k points (the general path cannot fill the machine there)
    assert cuda_lib.nnd_forward_kind(1, 8192, 8192) == 0 and cuda_lib.nnd_forward_kind(1, 10241, 100) == 1
    assert cuda_lib.nnd_forward_kind(2, 8192, 8192) == 2 and cuda_lib.nnd_workspace_bytes(1, 9000, 9000) == 0
    before = cuda_lib.nnd_set_small_max_points(0)
    try:
        assert cuda_lib.nnd_workspace_bytes(1, 256, 256) > 0 and cuda_lib.nnd_forward_kind(1, 256, 256) == 1
    finally:
        cuda_lib.nnd_set_small_max_points(before)
    assert cuda_lib.nnd_set_small_config(8, 3, 1) == 10001 and cuda_lib.nnd_set_small_config(8, 2, 3) == 10001
    assert cuda_lib.nnd_set_small_config(0, 0, 0) == 0


def test_pruned_search_workspace_and_validation_need_no_gpu(cuda_lib):
    # per cloud: two 8-byte key arrays + 16-byte sorted points (padded to 64-point tiles) + boxes; plus the sort's scratch
    w = cuda_lib.nnd_pruned_workspace_bytes(1, 1_000_000, 1_000_000)
    assert 2 * 1_000_000 * 32 <= w <= 2 * 1_000_000 * 32 + 32 * 1024 * 1024
    assert cuda_lib.nnd_pruned_workspace_bytes(0, 5, 5) == 256
    assert cuda_lib.nnd_pruned_workspace_bytes(70000, 40000, 10) == 0  # b * n beyond 2^31 - 1
    assert cuda_lib.nnd_forward_pruned_cuda_f32(None, None, -1, 1, 1, None, None, None, None, None, 0, None) == 10001
    assert cuda_lib.nnd_forward_pruned_cuda_f32(None, None, 0, 4, 4, None, None, None, None, None, 0, None) == 0
    assert cuda_lib.nnd_forward_pruned_cuda_f32(None, None, 1, 4, 4, None, None, None, None, None, 0, None) == 10001


def test_argument_validation_needs_no_gpu(cuda_lib):
    # negative sizes are rejected before any CUDA call
    assert cuda_lib.nnd_forward_cuda_f32(None, None, -1, 1, 1, None, None, None, None, None, 0, None) == 10001
    assert cuda_lib.nnd_backward_cuda_f32(None, None, None, None, None, None, 1, -5, 1, None, None, None) == 10001
    # b == 0: nothing to do, success without touching the device
    assert cuda_lib.nnd_forward_cuda_f32(None, None, 0, 4, 4, None, None, None, None, None, 0, None) == 0
    # row ranges must be whole row groups
    assert cuda_lib.nnd_forward_rows_cuda_f32(None, None, 1, 64, 64, 3, 64, None, None, None, None, 0, None) == 10001


def test_forward_kernel_sass_is_bit_exact_form():
    """The scan kernel must not contain a fused multiply-add of a real product: no scalar FFMA, and
    exactly 2 FFMA2 (the run-time 1.0 adds) per 3 FMUL2 / 3 FADD2 (grep the cubin)."""
    from pointcloudregistration_b200 import _cabi

    try:
        out = subprocess.run(["cuobjdump", "-sass", str(_cabi.LIB_PATH)], capture_output=True, text=True, check=True)
    except (FileNotFoundError, subprocess.CalledProcessError):
        pytest.skip("cuobjdump not available")
    kernels = re.split(r"Function : ", out.stdout)
    scan = [k for k in kernels if any(name in k.split("\n", 1)[0]
                                      for name in ("nnd_scan_kernel", "nnd_scan_sym_kernel", "nnd_scan_symc_kernel",
                                                   "nnd_small_kernel"))]
    assert len(scan) >= 3 + 3 + 6 + 6, "a forward kernel instantiation is missing from the cubin"
    for k in scan:
        ops = re.findall(r"^\s+/\*[0-9a-f]{4}\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_.]+)", k, flags=re.M)
        n = {name: sum(1 for o in ops if o.split(".")[0] == name) for name in ("FFMA", "FMUL2", "FADD2", "FFMA2", "FMNMX3")}
        assert n["FFMA"] == 0, f"scalar FFMA in the forward kernel: {n}"
        assert n["FMUL2"] > 0 and n["FMUL2"] == n["FADD2"], n
        assert 3 * n["FFMA2"] == 2 * n["FMUL2"], f"FFMA2:FMUL2 must be 2:3 (adds only), got {n}"
        assert n["FMNMX3"] > 0
    assert "UBLKCP" in out.stdout, "1-D TMA bulk copy (UBLKCP) missing from the forward kernel"


def test_panel_kernel_keeps_its_columns_in_uniform_registers():
    """The point of nnd_scan_symc_kernel: column points are loaded from the constant bank into UNIFORM
    registers (LDCU c[0x3][UR..]) and consumed as UR operands of the packed adds.  ptxas silently falls back to LDC
    into vector registers when the address arithmetic leaves the uniform datapath (tools/uniform_probe.cu); that
    costs ~7 % and no parity test would notice, so the cubin is checked."""
    from pointcloudregistration_b200 import _cabi

    try:
        out = subprocess.run(["cuobjdump", "-sass", str(_cabi.LIB_PATH)], capture_output=True, text=True, check=True)
    except (FileNotFoundError, subprocess.CalledProcessError):
        pytest.skip("cuobjdump not available")
    kernels = [k for k in re.split(r"Function : ", out.stdout) if "nnd_scan_symc_kernel" in k.split("\n", 1)[0]]
    assert len(kernels) == 6, "expected 3 panel sizes x (full, ragged) instantiations"
    for k in kernels:
        name = k.split("\n", 1)[0]
        assert len(re.findall(r"LDCU(?:\.64|\.128)? UR\d+, c\[0x3\]\[UR", k)) >= 40, f"{name}: panel not loaded by LDCU"
        assert not re.search(r"LDC(?:\.64|\.128)? R\d+, c\[0x3\]", k), f"{name}: panel loaded into vector registers"
        ur_adds = len(re.findall(r"FADD2 R\d+, -?R\d+(?:\.reuse)?\.F32, UR\d+\.F32x2", k))
        all_adds = len(re.findall(r"FADD2 ", k))
        assert ur_adds == all_adds > 0, f"{name}: {ur_adds} of {all_adds} FADD2 take the column pair from a uniform register"


def test_scan_kind_selection_needs_no_gpu(cuda_lib):
    """Which scan a row range takes (include/nnd_b200.h: nnd_rows_scan_kind): 1 persistent, 2 panel, 3 panel with
    the roles of the clouds swapped."""
    k = cuda_lib.nnd_rows_scan_kind
    assert cuda_lib.nnd_set_panel_min_rows(0) == 200000            # query only
    assert k(1, 1_000_000, 1_000_000, 0, 1_000_000) == 2              # C4, one GPU
    assert k(1, 1_000_000, 1_000_000, 125_000, 250_000) == 3          # C4, one of 8 ranks: shorter than xyz2 -> swapped
    assert k(1, 262_144, 262_144, 0, 262_144) == 2                    # C5
    assert k(1, 100_000, 100_000, 0, 100_000) == 1                    # C3
    assert k(1, 100_000, 1_000_000, 0, 100_000) == 3                  # N < M on one GPU
    assert k(1, 1_000_000, 100_000, 0, 1_000_000) == 2                # N > M: rows in registers as usual
    assert k(1, 16_384, 16_384, 0, 0) == 0
    # a long range of a cloud against a SMALL one through the rows ABI: many (row block x tile) units -> the large shape
    # -> panels, exactly what nnd_forward_rows_cuda_f32 does (the decision is one shared helper)
    assert k(1, 300_000, 4096, 0, 300_000) == 2
    before = cuda_lib.nnd_set_panel_min_rows(1)
    try:
        assert k(1, 9000, 8200, 0, 9000) == 2 and k(1, 9000, 30000, 0, 3000) == 3
    finally:
        cuda_lib.nnd_set_panel_min_rows(before)
    assert cuda_lib.nnd_set_panel_tiles(3) == 0 and cuda_lib.nnd_set_panel_tiles(-1) == 0   # invalid values only query
