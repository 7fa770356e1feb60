"""ctypes bindings of the two in-tree native libraries.

* ``libjordan_host.so``  — CPU-only: block-cyclic layout math, text I/O, CPU oracle.
* ``libjordan_b200.so``  — the sm_100a CUDA engine (contains the host objects as well).

The CUDA library is loaded lazily and *loudly*: on a machine with a GPU a missing or unloadable
extension raises instead of silently falling back to PyTorch.
"""
from __future__ import annotations

import ctypes as C
import os
from pathlib import Path

_LIBDIR = Path(__file__).resolve().parent / "lib"
_host = None
_cuda = None

JB_OK = 0
JB_ERR_CUDA = -1
JB_ERR_SINGULAR = -2
JB_ERR_NOMEM = -3
JB_ERR_ARG = -4
JB_ERR_IO = -5
JB_ERR_TIMEOUT = -6


class JbConfig(C.Structure):
    _fields_ = [
        ("n", C.c_int64),
        ("m", C.c_int32),
        ("p", C.c_int32),
        ("rank", C.c_int32),
        ("device", C.c_int32),
        ("keep_copy", C.c_int32),
        ("precision", C.c_int32),
    ]


class NativeError(RuntimeError):
    def __init__(self, code: int, msg: str):
        super().__init__(f"jordan_b200 error {code}: {msg}")
        self.code = code


class SingularMatrixError(NativeError):
    pass


def _declare_host(lib) -> None:
    i64, i32, dbl = C.c_int64, C.c_int32, C.c_double
    pd = C.POINTER(C.c_double)
    sig = {
        "jb_num_block_rows": (i64, [i64, i64]),
        "jb_find_sender": (i64, [i64, i64]),
        "jb_rows_per_process": (i64, [i64, i64, i64]),
        "jb_local_to_global_row": (i64, [i64, i64, i64, i64]),
        "jb_global_to_local_row": (i64, [i64, i64, i64]),
        "jb_owner_of_row": (i64, [i64, i64, i64]),
        "jb_padded_ld": (i64, [i64, i64]),
        "jb_formula": (dbl, [C.c_int, i64, i64, i64]),
        "jb_read_matrix_rows": (C.c_int, [C.c_char_p, i64, i64, i64, C.c_void_p]),
        "jb_read_matrix_full": (C.c_int, [C.c_char_p, i64, C.c_void_p]),
        "jb_matrix_reader_open": (C.c_int, [C.c_char_p, C.POINTER(C.c_void_p)]),
        "jb_matrix_reader_next": (C.c_int, [C.c_void_p, i64, C.c_void_p]),
        "jb_matrix_reader_close": (None, [C.c_void_p]),
        "jb_format_corner": (i64, [C.c_void_p, i64, i64, i64, C.c_char_p, i64]),
        "jb_cpu_block_jordan": (C.c_int, [C.c_void_p, C.c_void_p, i64, i64, i64, C.c_void_p]),
        "jb_cpu_inverse_block": (C.c_int, [C.c_void_p, C.c_void_p, dbl, i64]),
        "jb_cpu_minloc_fold": (C.c_int, [C.c_void_p, C.c_void_p, i64]),
        "jb_cpu_residual": (dbl, [C.c_void_p, C.c_void_p, i64]),
        "jb_plan_enumerate": (i64, [i32, i32, i32, i32, i64, i32, i32, i32, i32, i32, i32, i32, i32, i32, i32, i32,
                                    C.c_void_p, i64]),
    }
    for name, (res, args) in sig.items():
        fn = getattr(lib, name)
        fn.restype = res
        fn.argtypes = args
    del pd, i32


def _declare_cuda(lib) -> None:
    i64, i32 = C.c_int64, C.c_int32
    vp = C.c_void_p
    sig = {
        "jb_last_error": (C.c_char_p, []),
        "jb_engine_create": (C.c_int, [C.POINTER(JbConfig), C.POINTER(vp)]),
        "jb_engine_destroy": (None, [vp]),
        "jb_engine_symm_bytes": (C.c_size_t, [vp]),
        "jb_engine_alloc_symm": (C.c_int, [vp]),
        "jb_engine_bind_symm": (C.c_int, [vp, C.POINTER(vp), vp]),
        "jb_engine_ipc_handle": (C.c_int, [vp, C.c_int, vp]),
        "jb_engine_open_peers": (C.c_int, [vp, C.c_int, vp]),
        "jb_engine_reset_sync": (C.c_int, [vp]),
        "jb_engine_fill_formula": (C.c_int, [vp, C.c_int, vp]),
        "jb_engine_upload_rows": (C.c_int, [vp, vp, i64, i64, vp]),
        "jb_engine_finish_upload": (C.c_int, [vp, vp]),
        "jb_engine_local_norm": (C.c_int, [vp, C.POINTER(C.c_double)]),
        "jb_engine_set_norm": (C.c_int, [vp, C.c_double]),
        "jb_engine_invert": (C.c_int, [vp, vp]),
        "jb_engine_status": (C.c_int, [vp, vp]),
        "jb_engine_pivots": (C.c_int, [vp, vp]),
        "jb_engine_launch_count": (i64, [vp]),
        "jb_engine_set_profiling": (C.c_int, [vp, C.c_int]),
        "jb_symm_heap_create": (C.c_int, [C.c_size_t, C.c_int, C.c_int, C.c_int, C.c_int, C.POINTER(vp),
                                          C.POINTER(C.c_int)]),
        "jb_symm_heap_import": (C.c_int, [vp, C.POINTER(C.c_int)]),
        "jb_symm_heap_multicast_supported": (C.c_int, [vp]),
        "jb_symm_heap_mc_create": (C.c_int, [vp, C.POINTER(C.c_int)]),
        "jb_symm_heap_mc_import": (C.c_int, [vp, C.c_int]),
        "jb_symm_heap_mc_add": (C.c_int, [vp]),
        "jb_symm_heap_mc_bind": (C.c_int, [vp]),
        "jb_symm_heap_mc_drop": (None, [vp]),
        "jb_symm_heap_ptr": (vp, [vp, C.c_int]),
        "jb_symm_heap_mc_ptr": (vp, [vp]),
        "jb_symm_heap_bytes": (C.c_size_t, [vp]),
        "jb_symm_heap_destroy": (None, [vp]),
        "jb_engine_phase_times": (C.c_int, [vp, C.POINTER(C.c_double), C.POINTER(C.c_double)]),
        "jb_engine_w_ptr": (vp, [vp]),
        "jb_engine_ldw": (i64, [vp]),
        "jb_engine_local_rows": (i64, [vp]),
        "jb_engine_download": (C.c_int, [vp, C.c_int, i64, i64, i64, i64, vp, i64]),
        "jb_engine_residual": (C.c_int, [vp, C.c_int, C.POINTER(C.c_double), vp]),
        "jb_engine_refine_phase": (C.c_int, [vp, C.c_int, C.c_int, vp]),
        "jb_block_inverse": (C.c_int, [vp, i64, i32, C.c_double, vp, vp, vp, vp]),
        "jb_flag_litmus": (C.c_int, [C.c_int, C.c_int, C.c_int, C.POINTER(C.c_ulonglong)]),
        "jb_tc05_gemm_update": (C.c_int, [vp, i64, i64, i64, vp, i64, vp, i64, i32, i32, vp]),
    }
    for name, (res, args) in sig.items():
        if not hasattr(lib, name):
            continue
        fn = getattr(lib, name)
        fn.restype = res
        fn.argtypes = args


def _ensure_built(which: str) -> None:
    name = "libjordan_host.so" if which == "host" else "libjordan_b200.so"
    if (_LIBDIR / name).exists() and not os.environ.get("JB_REBUILD"):
        return
    from . import _build

    if which == "host":
        _build.build_host()
    else:
        _build.build_cuda()


def host():
    """The CPU-only library (always available)."""
    global _host
    if _host is None:
        _ensure_built("host")
        _host = C.CDLL(str(_LIBDIR / "libjordan_host.so"))
        _declare_host(_host)
    return _host


def cuda():
    """The CUDA engine.  Raises if it cannot be loaded (no silent fallback)."""
    global _cuda
    if _cuda is None:
        _ensure_built("cuda")
        path = Path(os.environ.get("JB_CUDA_LIB", str(_LIBDIR / "libjordan_b200.so")))
        try:
            _cuda = C.CDLL(str(path))
        except OSError as exc:  # pragma: no cover
            raise RuntimeError(f"native CUDA extension {path} failed to load: {exc}") from exc
        _declare_host(_cuda)
        _declare_cuda(_cuda)
    return _cuda


def cuda_loaded() -> bool:
    return _cuda is not None


def check(rc: int) -> None:
    if rc == JB_OK:
        return
    msg = cuda().jb_last_error().decode(errors="replace") if _cuda is not None else ""
    if rc == JB_ERR_SINGULAR:
        raise SingularMatrixError(rc, msg or "singular matrix")
    raise NativeError(rc, msg)
