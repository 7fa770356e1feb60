"""Python front end of the native symmetric heap (csrc/comm/symm_heap.cpp): one cuMemCreate
allocation per rank, descriptors exchanged between the rank processes (``Comm.exchange_fds``), every
peer mapped on every rank, plus the NVLS multicast alias where the GPUs sit behind an NVSwitch.
Replaces what the reference does with host buffers and MPI_Bcast / MPI_Send (main.cpp:1097,
1122-1129): the kernels store straight into the peers' copies."""
from __future__ import annotations

import ctypes as C
import os
from typing import Optional

from .. import _native
from .comm import Comm

_serial = 0


class NativeSymmHeap:
    """Collective over ``comm``.  ``ptrs[r]`` = rank r's region as mapped here, ``mc_ptr`` = multicast
    alias or 0.  ``NativeSymmHeap.create`` returns None (on every rank) if any rank cannot do it."""

    def __init__(self, handle, ptrs, mc_ptr, nbytes):
        self._h, self.ptrs, self.mc_ptr, self.nbytes = handle, ptrs, mc_ptr, nbytes

    @classmethod
    def create(cls, comm: Comm, nbytes: int, device: int, want_multicast: bool = True) -> Optional["NativeSymmHeap"]:
        global _serial
        lib, p = _native.cuda(), comm.world
        all_ok = lambda flag: comm.allreduce_sum_int(1 if flag else 0) == p  # noqa: E731
        heap, fd = C.c_void_p(), C.c_int(-1)
        rc = lib.jb_symm_heap_create(nbytes, device, comm.rank, p, int(want_multicast), C.byref(heap), C.byref(fd))
        if not all_ok(rc == 0):
            cls.last_error = lib.jb_last_error().decode(errors="replace")
            if rc == 0:
                os.close(fd.value)
                lib.jb_symm_heap_destroy(heap)
            return None
        _serial += 1
        fds = comm.exchange_fds(fd.value, tag=f"h{_serial}")
        os.close(fd.value)
        arr = (C.c_int * p)(*fds)
        arr[comm.rank] = -1
        rc = lib.jb_symm_heap_import(heap, arr)
        if not all_ok(rc == 0):
            cls.last_error = lib.jb_last_error().decode(errors="replace")
            lib.jb_symm_heap_destroy(heap)
            return None
        mc = all_ok(bool(lib.jb_symm_heap_multicast_supported(heap)))
        if mc:
            mfd = C.c_int(-1)
            ok0 = True
            if comm.rank == 0:
                ok0 = lib.jb_symm_heap_mc_create(heap, C.byref(mfd)) == 0
            if all_ok(ok0):
                got = comm.bcast_fd(mfd.value, src=0, tag=f"m{_serial}")
                ok = True
                if comm.rank == 0:
                    os.close(mfd.value)
                else:
                    ok = lib.jb_symm_heap_mc_import(heap, got) == 0
                ok = all_ok(ok)
                ok = ok and all_ok(lib.jb_symm_heap_mc_add(heap) == 0)     # vote = barrier: all added
                ok = ok and all_ok(lib.jb_symm_heap_mc_bind(heap) == 0)    # vote = barrier: all bound
                mc = ok
            else:
                mc = False
            if not mc:
                lib.jb_symm_heap_mc_drop(heap)
        ptrs = [int(lib.jb_symm_heap_ptr(heap, r) or 0) for r in range(p)]
        mcp = int(lib.jb_symm_heap_mc_ptr(heap) or 0) if mc else 0
        return cls(heap, ptrs, mcp, int(lib.jb_symm_heap_bytes(heap)))

    last_error = ""

    def close(self) -> None:
        if self._h:
            _native.cuda().jb_symm_heap_destroy(self._h)
            self._h = None
