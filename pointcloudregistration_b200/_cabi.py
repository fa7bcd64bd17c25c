"""ctypes binding of libnnd_b200.so (include/nnd_b200.h).

This is the only place Python touches native code on the product path.  If the
library is missing the import of any compute entry point raises: there is no CPU
or PyTorch fallback.
"""
from __future__ import annotations

import ctypes
from ctypes import c_char_p, c_float, c_int, c_longlong, c_size_t, c_void_p
from pathlib import Path

import os

# The product always loads the library that sits next to this file.  NND_B200_LIB (an experimental build of the same
# ABI, tools/variants.sh) is honoured ONLY together with NND_B200_DEBUG=1, and load() says so on stderr: a stray
# environment variable must not silently redirect the product to another library.
_override = os.environ.get("NND_B200_LIB") if os.environ.get("NND_B200_DEBUG") == "1" else None
LIB_PATH = Path(_override or Path(__file__).resolve().parent / "libnnd_b200.so")

#: every symbol include/nnd_b200.h declares -> (restype, argtypes)
SIGNATURES = {
    "nnd_version": (c_char_p, []),
    "nnd_error_string": (c_char_p, [c_int]),
    "nnd_workspace_bytes": (c_size_t, [c_int, c_int, c_int]),
    "nnd_query_workspace_bytes": (c_size_t, [c_int, c_int, c_int]),
    "nnd_forward_cuda_f32": (c_int, [c_void_p, c_void_p, c_int, c_int, c_int, c_void_p, c_void_p, c_void_p,
                                      c_void_p, c_void_p, c_size_t, c_void_p]),
    "nnd_forward_warm_cuda_f32": (c_int, [c_void_p, c_void_p, c_int, c_int, c_int, c_void_p, c_void_p, c_void_p,
                                           c_void_p, c_void_p, c_void_p, c_size_t, c_void_p]),
    "nnd_pruned_workspace_bytes": (c_size_t, [c_int, c_int, c_int]),
    "nnd_set_pruned_share": (c_int, [c_int]),
    "nnd_forward_pruned_reuse_cuda_f32": (c_int, [c_void_p, c_void_p, c_int, c_int, c_int, c_void_p, c_void_p, c_void_p,
                                                   c_void_p, c_void_p, c_size_t, c_int, c_void_p]),
    "nnd_forward_pruned_cuda_f32": (c_int, [c_void_p, c_void_p, c_int, c_int, c_int, c_void_p, c_void_p, c_void_p,
                                             c_void_p, c_void_p, c_size_t, c_void_p]),
    "nnd_query_pruned_cuda_f32": (c_int, [c_void_p, c_void_p, c_int, c_int, c_int, c_void_p, c_void_p, c_void_p,
                                           c_size_t, c_void_p]),
    "nnd_rows_workspace_bytes": (c_size_t, [c_int, c_int, c_int]),
    "nnd_colkeys_count": (c_size_t, [c_int, c_int]),
    "nnd_row_group": (c_int, []),
    "nnd_forward_is_symmetric": (c_int, [c_int, c_int]),
    "nnd_forward_rows_cuda_f32": (c_int, [c_void_p, c_void_p, c_int, c_int, c_int, c_int, c_int, c_void_p, c_void_p,
                                           c_void_p, c_void_p, c_size_t, c_void_p]),
    "nnd_finalize_cols_cuda_f32": (c_int, [c_void_p, c_void_p, c_int, c_int, c_int, c_void_p, c_void_p, c_void_p,
                                            c_void_p]),
    "nnd_peer_max_ranks": (c_int, []),
    "nnd_exchange_cols_p2p_cuda_f32": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_int, c_void_p,
                                                c_void_p, c_int, c_int, c_int, c_int, c_int, c_void_p]),
    "nnd_sum_p2p_cuda_f32": (c_int, [c_void_p, c_int, c_int, c_longlong, c_void_p]),
    "nnd_profile_enable": (c_int, [c_int]),
    "nnd_profile_collect": (c_int, [c_void_p, c_void_p]),
    "nnd_query_cuda_f32": (c_int, [c_void_p, c_void_p, c_int, c_int, c_int, c_void_p, c_void_p, c_void_p,
                                    c_size_t, c_void_p]),
    "nnd_query_noalloc_max_bytes": (c_size_t, []),
    "nnd_query_noalloc_cuda_f32": (c_int, [c_void_p, c_void_p, c_int, c_int, c_int, c_void_p, c_void_p, c_void_p]),
    "nnd_backward_cuda_f32": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_int,
                                       c_int, c_void_p, c_void_p, c_void_p]),
    "nnd_backward_term_cuda_f32": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_void_p,
                                            c_void_p, c_void_p]),
    "nnd_backward_exact_workspace_bytes": (c_size_t, [c_int, c_int, c_int]),
    "nnd_backward_exact_cuda_f32": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_int,
                                             c_int, c_void_p, c_void_p, c_void_p, c_size_t, c_void_p]),
    "nnd_last_launch_count": (c_int, []),
    "nnd_set_panel_min_rows": (c_longlong, [c_longlong]),
    "nnd_set_panel_tiles": (c_int, [c_int]),
    "nnd_rows_scan_kind": (c_int, [c_int, c_int, c_int, c_int, c_int]),
    "nnd_set_sym_seed_tiles": (c_int, [c_int]),
    "nnd_set_auto_hint": (c_int, [c_int]),
    "nnd_forward_kind": (c_int, [c_int, c_int, c_int]),
    "nnd_small_max_points": (c_int, []),
    "nnd_set_small_max_points": (c_int, [c_int]),
    "nnd_set_small_config": (c_int, [c_int, c_int, c_int]),
    "nnd_ragged_supported": (c_int, [c_int, c_int]),
    "nnd_forward_ragged_cuda_f32": (c_int, [c_void_p, c_void_p, c_int, c_int, c_int, c_void_p, c_void_p, c_void_p,
                                             c_void_p, c_void_p, c_void_p, c_void_p]),
    "nnd_query_ragged_cuda_f32": (c_int, [c_void_p, c_void_p, c_int, c_int, c_int, c_void_p, c_void_p, c_void_p,
                                           c_void_p, c_void_p]),
    "nnd_backward_ragged_cuda_f32": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_int,
                                              c_int, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p]),
    "nnd_knn_backward_ragged_cuda_f32": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_void_p,
                                                  c_void_p, c_void_p, c_void_p, c_void_p]),
    "nnd_chamfer_reduce_workspace_bytes": (c_size_t, []),
    "nnd_chamfer_reduce_cuda_f32": (c_int, [c_void_p, c_void_p, c_int, c_int, c_int, c_float, c_float, c_float, c_void_p,
                                             c_void_p, c_void_p, c_void_p, c_size_t, c_void_p]),
}

_lib = None


class NNDLibraryError(RuntimeError):
    pass


def load():
    """dlopen the library once and type its entry points.  Raises if it is not built."""
    global _lib
    if _lib is not None:
        return _lib
    if not LIB_PATH.exists():
        raise NNDLibraryError(
            f"{LIB_PATH} is missing. Build it with `python -c 'import __graft_entry__ as g; g.build()'` "
            "(nvcc, sm_100a). There is no CPU fallback for nnd.")
    if _override:
        import sys

        print(f"[nnd_b200] NND_B200_DEBUG=1: loading {LIB_PATH} instead of the packaged library", file=sys.stderr)
    lib = ctypes.CDLL(str(LIB_PATH))
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(lib, name)  # AttributeError if the header and the library disagree
        fn.restype = res
        fn.argtypes = args
    _lib = lib
    return lib


def check(code: int, what: str) -> None:
    if code != 0:
        msg = load().nnd_error_string(code).decode()
        raise RuntimeError(f"{what} failed: {msg} (code {code})")


def version() -> str:
    return load().nnd_version().decode()
