"""Flagship "model": distributed dense fp64 matrix inversion by block Gauss-Jordan elimination with
block-row pivoting, running on the hand-written sm_100a engine (``libjordan_b200.so``).

Public surface = what the reference program offers (``solve``/``Jordan``, ``main.cpp:343-519``,
``953-1204``): build or load ``A``, invert it with block size ``m`` over ``p`` ranks, report the
elapsed time of the elimination (max over ranks, ``main.cpp:450-458``), the top-left corner of the
inverse (``:461``) and the residual ``||A X - I||_inf`` (``:490-508``).
"""
from __future__ import annotations

import ctypes as C
import os
from typing import Optional

import torch

from .. import _native
from ..parallel.comm import Comm
from ..parallel.layout import Layout, global_to_local, owner_of_row

FORMULAS = {
    "absdiff": 1,   # |i-j|            main.cpp:52 (default)
    "hilbert": 2,   # 1/(i+j+1)        main.cpp:50 (-DHILBERT)
    "antidiag": 3,  # i+1 at j=n-1-i   main.cpp:53-56 (commented-out generator)
    "lehmer": 4,
    "diagdom": 5,
    "permuted": 6,
}


class _DevArray:
    """Zero-copy view of engine-owned device memory for ``torch.as_tensor``."""

    def __init__(self, ptr: int, shape, strides):
        self.__cuda_array_interface__ = {
            "shape": tuple(shape),
            "strides": tuple(strides),
            "typestr": "<f8",
            "data": (ptr, False),
            "version": 3,
        }


class JordanSolver:
    """One instance per rank (= per GPU).  All ranks must call the collective methods together."""

    def __init__(self, n: int, m: int, comm: Optional[Comm] = None, device: Optional[int] = None,
                 keep_copy: bool = False, symm: Optional[str] = None, precision: str = "fp64"):
        if not torch.cuda.is_available():
            raise RuntimeError("JordanSolver needs a CUDA device; use models.eager.EagerJordan on CPU")
        self.lib = _native.cuda()
        self.comm = comm or Comm()
        # the reference accepts any m, including m > n (one block row of height n, main.cpp:124-127,
        # 958): a block never needs to be larger than the matrix
        m = min(int(m), int(n))
        self.layout = Layout(n, m, self.comm.world, self.comm.rank)
        self.device = torch.cuda.current_device() if device is None else int(device)
        torch.cuda.set_device(self.device)
        self.formula: Optional[int] = None
        # precision: "fp64" = DMMA update (default); "tc05" = tcgen05 int8 error-free split update (K3b)
        self.precision = precision
        cfg = _native.JbConfig(n, m, self.comm.world, self.comm.rank, self.device, int(keep_copy),
                               {"fp64": 0, "tc05": 1}[precision])
        h = C.c_void_p()
        rc = self.lib.jb_engine_create(C.byref(cfg), C.byref(h))
        msg = self.lib.jb_last_error().decode(errors="replace") if rc != 0 else ""
        # collective vote (main.cpp:371-381): every rank learns the most severe code
        worst = -int(self.comm.allreduce_max(float(-rc)))
        if worst != 0:
            if rc == 0:
                self.lib.jb_engine_destroy(h)
            if worst == _native.JB_ERR_NOMEM:
                raise MemoryError("Not enough memory!")
            if worst == _native.JB_ERR_ARG:
                raise ValueError(msg or f"unsupported configuration n={n} m={m} (another rank rejected it)")
            raise _native.NativeError(worst, msg or "engine creation failed on another rank")
        self._h = h
        self._symm_tensor = None
        self._symm_handle = None
        self.multicast = False
        self._setup_symm(symm or os.environ.get("JB_SYMM", "auto"))
        self._open_w_peers()
        self.last_seconds = float("nan")

    # ------------------------------------------------------------------ set-up
    def _setup_symm(self, mode: str) -> None:
        """Symmetric (peer-visible) region.  ``auto``/``vmm``: the native heap (cuMemCreate, POSIX
        descriptors exchanged between the rank processes, NVLS multicast object;
        csrc/comm/symm_heap.cpp); ``torch``: torch.distributed._symmetric_memory; ``ipc``: cudaMalloc +
        cudaIpc handles (unicast only).  Every stage is voted on so that all ranks take one path."""
        lib, p = self.lib, self.comm.world
        self._heap = None
        if p == 1:
            _native.check(lib.jb_engine_alloc_symm(self._h))
            self.symm_mode = "local"
            return
        ok = False
        if mode in ("auto", "vmm"):
            ok = self._try_native_heap()
            if not ok and mode == "vmm":
                raise RuntimeError("native symmetric heap unavailable: " + getattr(self, "_symm_err", ""))
        if not ok and mode in ("auto", "torch"):
            ok = self._try_torch_symm()
            votes = self.comm.allreduce_sum_int(1 if ok else 0)
            if votes != p:
                ok = False
                self._symm_tensor = None
                self._symm_handle = None
                if mode == "torch":
                    raise RuntimeError("torch symmetric memory unavailable on some rank")
            if ok:
                hdl = self._symm_handle
                ptrs = (C.c_void_p * p)(*[C.c_void_p(int(x)) for x in hdl.buffer_ptrs])
                mc = int(getattr(hdl, "multicast_ptr", 0) or 0)
                if os.environ.get("JB_NO_MULTICAST"):
                    mc = 0
                _native.check(lib.jb_engine_bind_symm(self._h, ptrs, C.c_void_p(mc)))
                self.multicast = mc != 0
                self.symm_mode = "torch-symm" + ("+nvls" if mc else "")
        if not ok:
            _native.check(lib.jb_engine_alloc_symm(self._h))
            buf = C.create_string_buffer(64)
            _native.check(lib.jb_engine_ipc_handle(self._h, 0, buf))
            allh = b"".join(self.comm.allgather_bytes(buf.raw))
            _native.check(lib.jb_engine_open_peers(self._h, 0, allh))
            self.symm_mode = "cuda-ipc"
        torch.cuda.synchronize()
        self.comm.barrier()

    def _try_native_heap(self) -> bool:
        from ..parallel.symm import NativeSymmHeap

        p = self.comm.world
        heap = NativeSymmHeap.create(self.comm, int(self.lib.jb_engine_symm_bytes(self._h)), self.device,
                                     want_multicast=not os.environ.get("JB_NO_MULTICAST"))
        if heap is None:
            self._symm_err = NativeSymmHeap.last_error
            return False
        ptrs = (C.c_void_p * p)(*[C.c_void_p(x) for x in heap.ptrs])
        _native.check(self.lib.jb_engine_bind_symm(self._h, ptrs, C.c_void_p(heap.mc_ptr)))
        self._heap = heap
        self.multicast = heap.mc_ptr != 0
        self.symm_mode = "vmm" + ("+nvls" if heap.mc_ptr else "")
        return True

    def _try_torch_symm(self) -> bool:
        try:
            import torch.distributed as dist
            import torch.distributed._symmetric_memory as symm_mem

            nbytes = int(self.lib.jb_engine_symm_bytes(self._h))
            t = symm_mem.empty(nbytes, dtype=torch.uint8, device=torch.device("cuda", self.device))
            hdl = symm_mem.rendezvous(t, dist.group.WORLD)
            t.zero_()
            torch.cuda.synchronize()
            self._symm_tensor, self._symm_handle = t, hdl
            return True
        except Exception as exc:  # noqa: BLE001 - any failure selects the IPC path
            self._symm_err = repr(exc)
            return False

    def _open_w_peers(self) -> None:
        if self.comm.world == 1:
            return
        buf = C.create_string_buffer(64)
        _native.check(self.lib.jb_engine_ipc_handle(self._h, 1, buf))
        allh = b"".join(self.comm.allgather_bytes(buf.raw))
        _native.check(self.lib.jb_engine_open_peers(self._h, 1, allh))

    def _stream(self) -> C.c_void_p:
        return C.c_void_p(torch.cuda.current_stream().cuda_stream)

    def _agree_norm(self) -> float:
        loc = C.c_double()
        _native.check(self.lib.jb_engine_local_norm(self._h, C.byref(loc)))
        g = self.comm.allreduce_max(loc.value)  # global, not rank-local, ||A||inf
        _native.check(self.lib.jb_engine_set_norm(self._h, g))
        self.norm_a = g
        return g

    def fill_formula(self, formula) -> None:
        """K5: generate A from a formula id/name on the device (``init_matrix``, main.cpp:128-149)."""
        fid = FORMULAS[formula] if isinstance(formula, str) else int(formula)
        self.formula = fid
        _native.check(self.lib.jb_engine_fill_formula(self._h, fid, self._stream()))
        self._agree_norm()

    def upload_local_rows(self, rows: torch.Tensor) -> None:
        """Copy this rank's real rows (``rows_real x n``, CPU, ideally pinned) to the device."""
        lay = self.layout
        assert rows.dtype == torch.float64 and rows.shape == (lay.rows_real, lay.n), rows.shape
        rows = rows.contiguous()
        if lay.rows_real:
            _native.check(self.lib.jb_engine_upload_rows(self._h, C.c_void_p(rows.data_ptr()), 0,
                                                         lay.rows_real, self._stream()))
        _native.check(self.lib.jb_engine_finish_upload(self._h, self._stream()))
        self.formula = None
        self._agree_norm()

    # ------------------------------------------------------------------ the hot path
    def invert_async(self) -> None:
        """Enqueue all ``Nr`` elimination steps on the current stream (no host synchronisation)."""
        _native.check(self.lib.jb_engine_invert(self._h, self._stream()))

    def status(self) -> None:
        """Synchronise and agree on ONE verdict over the ranks (collective): raises
        :class:`SingularMatrixError` like the reference's rc -2, or ``NativeError`` with
        ``JB_ERR_TIMEOUT`` when a cross-GPU wait gave up on any rank.  Either way the step epochs of
        the ranks are inconsistent afterwards, so every rank resets its flags between two barriers
        and the engine stays usable."""
        rc = self.lib.jb_engine_status(self._h, self._stream())
        msg = self.lib.jb_last_error().decode(errors="replace") if rc != 0 else ""
        worst = -int(self.comm.allreduce_max(float(-rc)))  # also a barrier: every rank's stream is idle
        if worst in (_native.JB_ERR_SINGULAR, _native.JB_ERR_TIMEOUT):
            _native.check(self.lib.jb_engine_reset_sync(self._h))
            self.comm.barrier()
        if worst == _native.JB_ERR_SINGULAR:
            raise _native.SingularMatrixError(worst, msg or "singular matrix")
        if worst != 0:
            raise _native.NativeError(worst, msg or "a peer rank reported an error")

    def invert(self) -> float:
        """Timed inversion: device time of the elimination, max over ranks (main.cpp:427-455)."""
        torch.cuda.synchronize()
        self.comm.barrier()
        e0 = torch.cuda.Event(enable_timing=True)
        e1 = torch.cuda.Event(enable_timing=True)
        e0.record()
        self.invert_async()
        e1.record()
        self.status()
        self.last_seconds = self.comm.allreduce_max(e0.elapsed_time(e1) * 1e-3)
        self.comm.barrier()
        return self.last_seconds

    # ------------------------------------------------------------------ results
    def local_inverse(self) -> torch.Tensor:
        """Zero-copy ``rows_real x n`` view of this rank's rows of ``X = A^-1``."""
        lay = self.layout
        ptr = int(self.lib.jb_engine_w_ptr(self._h))
        ldw = int(self.lib.jb_engine_ldw(self._h))
        if lay.rows_real == 0:
            return torch.empty(0, lay.n, dtype=torch.float64, device=f"cuda:{self.device}")
        arr = _DevArray(ptr, (lay.rows_real, lay.n), (ldw * 8, 8))
        return torch.as_tensor(arr, device=f"cuda:{self.device}")

    def pivots(self) -> list[int]:
        buf = (C.c_int32 * self.layout.Nr)()
        _native.check(self.lib.jb_engine_pivots(self._h, buf))
        return list(buf)

    def gather_inverse(self) -> Optional[torch.Tensor]:
        """Full inverse on rank 0 (small n only; test / tooling helper)."""
        lay = self.layout
        loc = self.local_inverse().cpu()
        parts = self.comm.allgather_obj((lay.global_rows(), loc.numpy()))
        if self.comm.rank != 0:
            return None
        out = torch.empty(lay.n, lay.n, dtype=torch.float64)
        for rows, data in parts:
            if len(rows):
                out[torch.tensor(rows)] = torch.from_numpy(data)
        return out

    def corner(self, which: str = "X", k: int = 10) -> Optional[torch.Tensor]:
        """Top-left ``min(n,k)^2`` corner gathered on rank 0 (``print_matrix``, main.cpp:297-341)."""
        lay = self.layout
        nm = min(lay.n, k)
        sel = 0 if which == "X" else 1
        mine = []
        for gi in range(nm):
            if owner_of_row(gi, lay.m, lay.p) == lay.rank:
                row = torch.empty(nm, dtype=torch.float64)
                _native.check(self.lib.jb_engine_download(self._h, sel, global_to_local(gi, lay.m, lay.p),
                                                          1, 0, nm, C.c_void_p(row.data_ptr()), nm))
                mine.append((gi, row))
        parts = self.comm.allgather_obj(mine)
        if self.comm.rank != 0:
            return None
        out = torch.zeros(nm, nm, dtype=torch.float64)
        for part in parts:
            for gi, row in part:
                out[gi] = row
        return out

    def residual(self, formula=None) -> float:
        """``||A X - I||_inf`` with the ORIGINAL A (regenerated from the formula, or the kept copy)."""
        fid = self.formula if formula is None else (FORMULAS[formula] if isinstance(formula, str) else formula)
        if fid is None:
            fid = -1
        torch.cuda.synchronize()
        self.comm.barrier()  # every rank's rows of X are final before anybody pulls them
        loc = C.c_double()
        _native.check(self.lib.jb_engine_residual(self._h, int(fid), C.byref(loc), self._stream()))
        self.comm.barrier()
        return self.comm.allreduce_max(loc.value)

    def refine(self, steps: int = 1) -> None:
        """K7 Newton-Schulz refinement ``X <- X + X (I - A X)`` in fp64, collective over all ranks."""
        fid = -1 if self.formula is None else self.formula
        if self.comm.world > 1 and not getattr(self, "_r_peers_open", False):
            buf = C.create_string_buffer(64)
            _native.check(self.lib.jb_engine_ipc_handle(self._h, 2, buf))
            allh = b"".join(self.comm.allgather_bytes(buf.raw))
            _native.check(self.lib.jb_engine_open_peers(self._h, 2, allh))
            self._r_peers_open = True
        for _ in range(steps):
            torch.cuda.synchronize()
            self.comm.barrier()  # X is final everywhere
            _native.check(self.lib.jb_engine_refine_phase(self._h, fid, 0, self._stream()))
            torch.cuda.synchronize()
            self.comm.barrier()  # every residual slab is complete before anybody pulls it
            _native.check(self.lib.jb_engine_refine_phase(self._h, fid, 1, self._stream()))
            torch.cuda.synchronize()
            self.comm.barrier()

    def set_profiling(self, on: bool = True) -> None:
        _native.check(self.lib.jb_engine_set_profiling(self._h, int(on)))

    def phase_times(self) -> dict:
        """Device time (ms) of the last inversion, split by schedule segment and summed over the
        steps.  Needs ``set_profiling(True)``."""
        seg, tail = (C.c_double * 4)(), C.c_double()
        _native.check(self.lib.jb_engine_phase_times(self._h, seg, C.byref(tail)))
        if os.environ.get("JB_NO_LOOKAHEAD") or self.layout.Nr == 1:
            # serial schedule: fused step kernel of step t, then K1(t+1)
            return {"step_ms": seg[0], "k1_ms": seg[2], "tail_ms": tail.value, "lookahead": False}
        if self.precision == "tc05":
            return {"scale_swap_ms": seg[0], "tc05_update_ms": seg[1], "k1_ms": seg[2], "tail_ms": tail.value,
                    "lookahead": False}
        # look-ahead: kernel durations (they overlap, so they do not add up to the inversion time) of
        # A(t) (swap + block columns t, t+1; beside B2(t-1)), B1(t), B2(t) (next pivot row scaled +
        # broadcast, rest of the update) and K1(t+1) incl. its wait for SMs
        return {"stepA_ms": seg[0], "stepB1_ms": seg[1], "stepB2_ms": seg[2], "k1_ms": seg[3],
                "tail_ms": tail.value, "lookahead": True}

    @property
    def launch_count(self) -> int:
        return int(self.lib.jb_engine_launch_count(self._h))

    def close(self) -> None:
        if getattr(self, "_h", None):
            torch.cuda.synchronize()
            self.comm.barrier()
            self.lib.jb_engine_destroy(self._h)
            self._h = None
            if getattr(self, "_heap", None):
                self._heap.close()   # the heap outlives the engine bound to it
                self._heap = None

    def __del__(self):  # pragma: no cover
        try:
            if getattr(self, "_h", None):
                self.lib.jb_engine_destroy(self._h)
                self._h = None
        except Exception:
            pass


def invert(a: torch.Tensor, m: int = 128) -> torch.Tensor:
    """Single-GPU convenience: invert a dense fp64 matrix held on the host or the device."""
    assert a.dim() == 2 and a.shape[0] == a.shape[1]
    n = a.shape[0]
    s = JordanSolver(n, m, comm=_SingleComm())
    try:
        s.upload_local_rows(a.detach().to("cpu", torch.float64))
        s.invert()
        return s.local_inverse().clone()
    finally:
        s.close()


class _SingleComm(Comm):
    """A world of one, regardless of any initialised process group."""

    def __init__(self):
        self.enabled, self.rank, self.world, self.backend, self.ctrl = False, 0, 1, "none", None
