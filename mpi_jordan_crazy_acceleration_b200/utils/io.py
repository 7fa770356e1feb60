"""Matrix text I/O: the reference's input format and its scatter / gather pattern.

``read_matrix`` (``main.cpp:209-282``): the file holds ``n*n`` whitespace-separated numbers,
row-major, parsed with ``%lf``; ONE rank reads it a block row at a time and sends each block row to
its owner (C4).  ``print_matrix`` (``main.cpp:297-341``) gathers the top-left ``min(n,10)^2`` corner on
rank 0 (C5) and prints ``"%.2f\\t"`` per value.  Parsing is native (``csrc/host/io_text.cpp``).
"""
from __future__ import annotations

import ctypes as C
from typing import Optional

import torch

from .. import _native
from ..parallel.comm import Comm
from ..parallel.layout import Layout

MAX_P = 10  # print size (main.cpp:6)


class MatrixFileError(RuntimeError):
    """code -1: cannot open, -2: cannot read (main.cpp:392-394)."""

    def __init__(self, code: int, path: str):
        self.code = code
        self.path = path
        super().__init__(("cannot open %s" if code == -1 else "cannot read %s") % path)


def read_full(path: str, n: int) -> torch.Tensor:
    out = torch.empty(n, n, dtype=torch.float64)
    rc = _native.host().jb_read_matrix_full(path.encode(), n, C.c_void_p(out.data_ptr()))
    if rc:
        raise MatrixFileError(rc, path)
    return out


def write_matrix(path: str, a: torch.Tensor) -> None:
    with open(path, "w") as f:
        for row in a.tolist():
            f.write(" ".join(repr(float(v)) for v in row) + "\n")


def scatter_text_matrix(path: str, layout: Layout, comm: Comm, reader: int = 0,
                        pin: bool = False) -> torch.Tensor:
    """Returns this rank's ``rows_real x n`` rows.  The reading rank scans the file ONCE and holds a
    single block row (``m*n`` numbers) at a time, which it keeps or sends to the owner - the
    reference's streaming scatter (``main.cpp:249-263``), not a full ``n*n`` staging copy.  All ranks
    agree on errors (C3); a short or garbled file stops the scatter at the block row where it is
    detected (fail fast)."""
    n, m, p = layout.n, layout.m, layout.p
    lib = _native.host()
    local = torch.empty(layout.rows_real, n, dtype=torch.float64)
    if pin and torch.cuda.is_available():
        local = local.pin_memory()
    rd = C.c_void_p()
    err = 0
    if comm.rank == reader:
        err = lib.jb_matrix_reader_open(path.encode(), C.byref(rd))
    err = comm.bcast_obj(err, src=reader)
    if err:
        raise MatrixFileError(err, path)
    blk = torch.empty(m, n, dtype=torch.float64)
    try:
        for I in range(layout.Nr):
            owner = I % p
            h = min(m, n - I * m)
            slot = I // p
            if comm.rank == reader:
                err = lib.jb_matrix_reader_next(rd, h * n, C.c_void_p(blk.data_ptr()))
            if comm.world > 1:
                err = comm.bcast_obj(err, src=reader)
            if err:
                raise MatrixFileError(err, path)
            if comm.rank == reader:
                if owner == reader:
                    local[slot * m: slot * m + h] = blk[:h]
                else:
                    comm.send_tensor(blk[:h].contiguous(), owner)
            elif comm.rank == owner:
                buf = torch.empty(h, n, dtype=torch.float64)
                comm.recv_tensor(buf, reader)
                local[slot * m: slot * m + h] = buf
    finally:
        if comm.rank == reader and rd:
            lib.jb_matrix_reader_close(rd)
    return local


def format_corner(a: torch.Tensor) -> str:
    """``print_row`` formatting (main.cpp:284-295)."""
    a = a.contiguous().to(torch.float64)
    rows, cols = a.shape
    buf = C.create_string_buffer(rows * cols * 32 + 64)
    w = _native.host().jb_format_corner(C.c_void_p(a.data_ptr()), rows, cols, cols, buf, len(buf))
    if w < 0:
        raise RuntimeError("corner buffer too small")
    return buf.value.decode()


def corner_from_rows(local_rows: torch.Tensor, layout: Layout, comm: Comm, k: int = MAX_P) -> Optional[torch.Tensor]:
    nm = min(layout.n, k)
    mine = []
    for lr, gi in enumerate(layout.global_rows()):
        if gi < nm:
            mine.append((gi, local_rows[lr, :nm].cpu().clone()))
    parts = comm.allgather_obj(mine)
    if comm.rank != 0:
        return None
    out = torch.zeros(nm, nm, dtype=torch.float64)
    for part in parts:
        for gi, row in part:
            out[gi] = row
    return out
