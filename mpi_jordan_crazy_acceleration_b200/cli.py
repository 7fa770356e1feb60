"""Command line with the reference's contract (``main``/``solve``, main.cpp:65-93, 343-519).

    python -m mpi_jordan_crazy_acceleration_b200.cli n m [<file>]
    python -m mpi_jordan_crazy_acceleration_b200.cli n m r s [<file>]     # print size r, formula id s
    ... --print r --formula s --backend {auto,engine,eager}

Under ``torchrun`` every process is one rank (= one GPU).  Rank 0 prints the report:
``A`` + corner, ``glob_time: %.2f``, ``inverse matrix:`` + blank + corner, ``residual: %e``.
Exit codes 0 / 1 (usage) / 2 (cannot open, cannot read, singular matrix, out of memory).
On a machine without a GPU the same report is produced by the library-call path (EagerJordan).
"""
from __future__ import annotations

import sys
import time

import torch

from . import _native
from .parallel.comm import Comm
from .parallel.layout import Layout
from .utils.io import MAX_P, MatrixFileError, corner_from_rows, format_corner, scatter_text_matrix
from .utils.report import report_header, report_residual, report_time, usage_line


def _parse(argv):
    print_size, formula, backend = MAX_P, 1, "auto"
    pos = []
    i = 0
    bad = False
    while i < len(argv):
        a = argv[i]
        if a == "--print" and i + 1 < len(argv):
            print_size = _atoi(argv[i + 1]); i += 2
        elif a == "--formula" and i + 1 < len(argv):
            formula = _atoi(argv[i + 1]); i += 2
        elif a == "--backend" and i + 1 < len(argv):
            backend = argv[i + 1]; i += 2
        elif a.startswith("--"):
            bad = True; i += 1
        else:
            pos.append(a); i += 1
    name = None
    n = m = 0
    if len(pos) >= 2:
        n, m = _atoi(pos[0]), _atoi(pos[1])
    if len(pos) == 3:
        name = pos[2]
    elif len(pos) in (4, 5):
        print_size, formula = _atoi(pos[2]), _atoi(pos[3])
        if len(pos) == 5:
            name = pos[4]
    ok = (not bad and 2 <= len(pos) <= 5 and n > 0 and m > 0 and print_size > 0 and formula >= 0
          and not (formula == 0 and name is None) and backend in ("auto", "engine", "eager"))
    return ok, n, m, print_size, formula, name, backend


def _atoi(s: str) -> int:
    """C atoi: leading integer, 0 on garbage (main.cpp:77 treats 0 as a usage error)."""
    s = s.strip()
    k = 0
    if k < len(s) and s[k] in "+-":
        k += 1
    while k < len(s) and s[k].isdigit():
        k += 1
    try:
        return int(s[:k])
    except ValueError:
        return 0


def main(argv=None, out=sys.stdout) -> int:
    argv = list(sys.argv[1:] if argv is None else argv)
    comm = Comm.init_from_env()
    ok, n, m, print_size, formula, name, backend = _parse(argv)
    say = (lambda s: (out.write(s), out.flush())) if comm.rank == 0 else (lambda s: None)
    if not ok:
        say(usage_line(sys.argv[0] if argv is not None else "jordan"))
        return 1
    use_engine = backend == "engine" or (backend == "auto" and torch.cuda.is_available())
    lay = Layout(n, m, comm.world, comm.rank)
    try:
        if use_engine:
            from .models.jordan import JordanSolver
            solver = JordanSolver(n, m, comm=comm, keep_copy=name is not None)
        else:
            from .models.eager import EagerJordan
            solver = EagerJordan(n, m, comm=comm)
    except MemoryError:
        say("Not enough memory!\n")
        return 2
    try:
        if name is not None:
            try:
                rows = scatter_text_matrix(name, lay, comm, pin=use_engine)
            except MatrixFileError as exc:
                say(str(exc) + "\n")
                return 2
            solver.upload_local_rows(rows)
            a_rows = rows
        else:
            solver.fill_formula(formula)
            a_rows = None
        say(report_header())
        if use_engine:
            ca = solver.corner("A" if name is not None else "X", print_size)  # before inversion W holds A
        else:
            ca = corner_from_rows(solver.W[: lay.rows_real] if a_rows is None else a_rows, lay, comm, print_size)
        if comm.rank == 0:
            say(format_corner(ca))
        try:
            if use_engine:
                seconds = solver.invert()
            else:
                comm.barrier()
                t0 = time.perf_counter()
                solver.invert()
                seconds = comm.allreduce_max(time.perf_counter() - t0)
        except _native.SingularMatrixError:
            say("singular matrix\n")
            return 2
        say(report_time(seconds))
        cx = solver.corner("X", print_size) if use_engine else corner_from_rows(solver.local_inverse(), lay, comm, print_size)
        if comm.rank == 0:
            say(format_corner(cx))
        say(report_residual(solver.residual()))
        return 0
    finally:
        if use_engine:
            solver.close()


if __name__ == "__main__":
    rc = main()
    try:
        import torch.distributed as dist

        if dist.is_initialized():
            dist.destroy_process_group()
    except Exception:  # noqa: BLE001
        pass
    sys.exit(rc)
