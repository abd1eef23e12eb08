"""ctypes binding of libbibfs_b200.so (C ABI in include/bibfs_b200.h).

The shared library is built in-tree (``make`` or ``__graft_entry__.build()``) into
``bidirectional_bfs_b200/_native/`` so that it travels with the repository snapshot to the GPU
box.  Nothing here falls back to a Python implementation: if the library is missing it is built,
and if that fails the import error says so.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess
import threading
from pathlib import Path

_PKG = Path(__file__).resolve().parent
_ROOT = _PKG.parent
_SO = Path(os.environ["BIBFS_B200_LIB"]) if os.environ.get("BIBFS_B200_LIB") else _PKG / "_native" / "libbibfs_b200.so"
_lock = threading.Lock()
_lib = None

MAX_RANKS = 8
NO_VERTEX = 0xFFFFFFFF

# query flags (bibfs_b200.h)
F_TRACE, F_FORCE_GRID, F_FORCE_TD, F_FORCE_BU, F_NO_PATH, F_DEVICE_CLOCK, F_NO_FRONT = 1, 2, 4, 8, 16, 32, 64
STATUS_HANDED_OVER = 100


class GenSpec(C.Structure):
    _fields_ = [
        ("kind", C.c_int32), ("scale", C.c_uint32), ("n", C.c_uint64), ("m", C.c_uint64),
        ("p", C.c_double), ("a16", C.c_uint32), ("ab16", C.c_uint32), ("abc16", C.c_uint32),
        ("width", C.c_uint32), ("height", C.c_uint32), ("seed", C.c_uint64),
    ]


class Result(C.Structure):
    _fields_ = [
        ("length", C.c_int32), ("meet", C.c_uint32), ("depth_s", C.c_uint32), ("depth_t", C.c_uint32),
        ("levels", C.c_uint32), ("levels_bottom_up", C.c_uint32), ("path_len", C.c_uint32),
        ("status", C.c_int32), ("edges_scanned", C.c_uint64), ("vertices_visited", C.c_uint64),
        ("search_ms", C.c_double), ("path_ms", C.c_double),
        ("levels_grid", C.c_uint32), ("levels_paired", C.c_uint32),
    ]


class CtxDesc(C.Structure):
    _fields_ = [
        ("n", C.c_uint32), ("nranks", C.c_uint32), ("rank_lo", C.c_uint32), ("vranks", C.c_uint32),
        ("d_row_ptr", C.c_void_p * MAX_RANKS), ("d_col", C.c_void_p * MAX_RANKS),
        ("max_path", C.c_uint32), ("solo_threshold", C.c_uint32), ("bottom_up_div", C.c_uint32),
        ("max_ctas", C.c_uint32), ("watchdog_ms", C.c_uint32), ("max_degree", C.c_uint32),
        ("d_vrec", C.c_void_p * MAX_RANKS),
    ]


class LevelTrace(C.Structure):
    _fields_ = [
        ("side", C.c_uint32), ("mode", C.c_uint32), ("frontier", C.c_uint32), ("discovered", C.c_uint32),
        ("t_ns", C.c_uint64), ("edges", C.c_uint64), ("t_expand_ns", C.c_uint64),
    ]


def build_native(verbose: bool = False) -> Path:
    """Compile the library for sm_100a with nvcc (cross-compiles without a GPU)."""
    cmd = ["make", "-C", str(_ROOT), "-j", str(os.cpu_count() or 4)]
    proc = subprocess.run(cmd, capture_output=True, text=True)
    if verbose or proc.returncode != 0:
        print(proc.stdout[-4000:])
        print(proc.stderr[-4000:])
    if proc.returncode != 0 or not _SO.exists():
        raise RuntimeError("building libbibfs_b200.so failed (see output above)")
    return _SO


def _declare(lib):
    vp, u32, u64, i32 = C.c_void_p, C.c_uint32, C.c_uint64, C.c_int32
    P = C.POINTER
    sig = {
        "bibfs_last_error": (C.c_char_p, []),
        "bibfs_version": (C.c_char_p, []),
        "bibfs_launch_count": (u64, []),
        "bibfs_bin_header": (C.c_int, [C.c_char_p, P(u32), P(u64)]),
        "bibfs_bin_read": (C.c_int, [C.c_char_p, vp, u64]),
        "bibfs_bin_write": (C.c_int, [C.c_char_p, u32, u64, vp]),
        "bibfs_dense_header": (C.c_int, [C.c_char_p, P(u32)]),
        "bibfs_dense_read": (C.c_int, [C.c_char_p, vp, u32]),
        "bibfs_gen_count_host": (C.c_int, [P(GenSpec), P(u64)]),
        "bibfs_gen_edges_host": (C.c_int, [P(GenSpec), vp, u64, P(u64)]),
        "bibfs_gen_workspace_bytes": (C.c_size_t, [P(GenSpec)]),
        "bibfs_gen_count_device": (C.c_int, [P(GenSpec), vp, C.c_size_t, P(u64), vp]),
        "bibfs_gen_edges_device": (C.c_int, [P(GenSpec), vp, u64, vp, C.c_size_t, P(u64), vp]),
        "bibfs_csr_count_host": (C.c_int, [u32, u64, vp, P(GenSpec), u32, u32, vp]),
        "bibfs_csr_fill_host": (C.c_int, [u32, u64, vp, P(GenSpec), u32, u32, vp, vp, C.c_int]),
        "bibfs_csr_workspace_bytes": (C.c_size_t, [u32]),
        "bibfs_csr_count_device": (C.c_int, [u32, u64, vp, P(GenSpec), u32, u32, vp, vp, C.c_size_t, P(u64), vp]),
        "bibfs_csr_fill_device": (C.c_int, [u32, u64, vp, P(GenSpec), u32, u32, vp, vp, vp, C.c_size_t, C.c_int, vp]),
        "bibfs_scan_workspace_bytes": (C.c_size_t, [u64]),
        "bibfs_scan_u32_device": (C.c_int, [vp, vp, u64, vp, C.c_size_t, vp]),
        "bibfs_scan_noalloc_max_count": (u64, []),
        "bibfs_scan_u32_device_noalloc": (C.c_int, [vp, vp, u64, vp]),
        "bibfs_serial": (C.c_int, [u32, vp, vp, u32, u32, C.c_int, P(Result), vp, u32]),
        "bibfs_cpu_ranks": (C.c_int, [u32, vp, vp, u32, u32, C.c_int, P(Result), vp, u32]),
        "bibfs_validate_path": (C.c_int, [u32, vp, vp, u32, u32, vp, u32]),
        "bibfs_block_size": (u32, [u32, u32]),
        "bibfs_gen_chunks": (u64, [P(GenSpec)]),
        "bibfs_csr_route_degree_device": (C.c_int, [u32, u64, vp, P(GenSpec), u64, u64, u32, vp, vp, vp]),
        "bibfs_csr_route_fill_device": (C.c_int, [u32, u64, vp, P(GenSpec), u64, u64, u32, vp, vp, vp]),
        "bibfs_csr_sort_rows_device": (C.c_int, [vp, vp, u32, vp]),
        "bibfs_vrec_bytes": (C.c_size_t, [u32]),
        "bibfs_vrec_build_device": (C.c_int, [vp, vp, u32, u64, vp, vp, vp]),
        "bibfs_ctx_create": (C.c_int, [P(vp), P(CtxDesc)]),
        "bibfs_ctx_destroy": (C.c_int, [vp]),
        "bibfs_ctx_ipc_handle": (C.c_int, [vp, u32, vp]),
        "bibfs_ctx_attach_ipc": (C.c_int, [vp, u32, vp]),
        "bibfs_ctx_attach_local": (C.c_int, [vp, u32, vp]),
        "bibfs_ctx_arena_bytes": (C.c_size_t, [vp]),
        "bibfs_ctx_reset": (C.c_int, [vp, vp]),
        "bibfs_ctx_set_peer_graph": (C.c_int, [vp, u32, vp, vp]),
        "bibfs_ctx_resume": (C.c_int, [vp, vp]),
        "bibfs_ipc_export": (C.c_int, [vp, vp, P(u64)]),
        "bibfs_ipc_import": (C.c_int, [vp, u64, P(vp)]),
        "bibfs_copy_device": (C.c_int, [vp, vp, C.c_size_t, vp]),
        "bibfs_ctx_query": (C.c_int, [vp, u32, u32, vp, u32, vp]),
        "bibfs_ctx_result": (C.c_int, [vp, P(Result), vp, u32, vp]),
        "bibfs_ctx_query_result": (C.c_int, [vp, u32, u32, vp, u32, P(Result), vp, u32, vp]),
        "bibfs_ctx_timing": (C.c_int, [vp, P(C.c_float), P(C.c_float)]),
        "bibfs_ctx_trace": (C.c_int, [vp, u32, P(LevelTrace), u32, P(u32)]),
        "bibfs_batch_workspace_bytes": (C.c_size_t, [u32]),
        "bibfs_batch_query_device": (C.c_int, [u32, vp, vp, vp, vp, u32, vp, vp, vp, C.c_size_t, P(C.c_float), vp]),
        "bibfs_pool_create": (C.c_int, [P(vp), P(CtxDesc), u32, u32]),
        "bibfs_pool_query": (C.c_int, [vp, vp, vp, u32, vp, vp, vp, vp, P(C.c_float)]),
        "bibfs_pool_destroy": (C.c_int, [vp]),
        "bibfs_pool_set_peer_graph": (C.c_int, [vp, u32, vp, vp]),
        "bibfs_resident_start": (C.c_int, [P(vp), P(CtxDesc), u32, u32]),
        "bibfs_resident_query": (C.c_int, [vp, u32, u32, u32, P(Result), vp, u32]),
        "bibfs_resident_launches": (C.c_int, [vp, P(C.c_uint64)]),
        "bibfs_resident_stop": (C.c_int, [vp]),
        "bibfs_naive_levelsync_device": (C.c_int, [u32, vp, vp, u32, u32, P(i32), P(C.c_float), vp]),
    }
    for name, (res, args) in sig.items():
        fn = getattr(lib, name)
        fn.restype = res
        fn.argtypes = args


def lib():
    """The loaded library (built on first use if the .so is missing)."""
    global _lib
    if _lib is None:
        with _lock:
            if _lib is None:
                if not _SO.exists():
                    build_native()
                handle = C.CDLL(str(_SO), mode=C.RTLD_GLOBAL)
                _declare(handle)
                _lib = handle
    return _lib


class BibfsError(RuntimeError):
    pass


def check(rc: int) -> None:
    if rc != 0:
        raise BibfsError(lib().bibfs_last_error().decode("utf-8", "replace"))


def native_path() -> Path:
    return _SO
