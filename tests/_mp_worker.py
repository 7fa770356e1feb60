"""torchrun worker: multi-GPU correctness of the fused engine (peer mailboxes, broadcast, swap).
usage: torchrun --nproc-per-node P tests/_mp_worker.py [symm-mode]"""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import torch  # noqa: E402

from mpi_jordan_crazy_acceleration_b200 import _native  # noqa: E402
from mpi_jordan_crazy_acceleration_b200.models.eager import formula_rows  # noqa: E402
from mpi_jordan_crazy_acceleration_b200.models.jordan import JordanSolver  # noqa: E402
from mpi_jordan_crazy_acceleration_b200.ops import cpu_block_jordan  # noqa: E402
from mpi_jordan_crazy_acceleration_b200.parallel.comm import Comm  # noqa: E402


def main():
    mode = sys.argv[1] if len(sys.argv) > 1 else "auto"
    comm = Comm.init_from_env()
    dev = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(dev)
    p = comm.world
    cases = [(1, 512, 64), (6, 512, 64), (3, 96, 8), (1, 300, 7), (6, 1024, 128), (1, 768, 256),
             (5, 500, 60), (1, 30, 8), (6, 2048, 32), (1, 2048, 128)]
    results = []
    for fid, n, m in cases:
        s = JordanSolver(n, m, comm=comm, device=dev, symm=mode)
        for rep in range(2):
            s.fill_formula(fid)
            sec = s.invert()
        X = s.gather_inverse()
        piv = s.pivots()
        res = s.residual()
        rec = {"fid": fid, "n": n, "m": m, "p": p, "sec": sec, "residual": res, "symm": s.symm_mode}
        if comm.rank == 0:
            A = formula_rows(fid, n, torch.arange(n))
            Xo, pivo = cpu_block_jordan(A, m, p)
            rec["piv_ok"] = piv == pivo
            rec["maxerr"] = float((X - Xo).abs().max() / Xo.abs().max())
            rec["swaps"] = sum(1 for t, g in enumerate(pivo) if g != t)
            rec["ok"] = bool(rec["piv_ok"] and rec["maxerr"] < 1e-8)
        results.append(rec)
        s.close()
    # singular verdict must be unanimous and recoverable
    s = JordanSolver(64, 16, comm=comm, device=dev, symm=mode)
    s.fill_formula(2) if False else None
    lay = s.layout
    s.upload_local_rows(torch.ones(lay.rows_real, 64, dtype=torch.float64))
    sing = False
    try:
        s.invert()
    except _native.SingularMatrixError:
        sing = True
    s.fill_formula(1)
    s.invert()
    X = s.gather_inverse()
    if comm.rank == 0:
        ref = torch.linalg.inv(formula_rows(1, 64, torch.arange(64)))
        results.append({"case": "singular+recover", "ok": bool(sing and (X - ref).abs().max() < 1e-9)})
    s.close()
    # K7 refinement across ranks: damage X, one Newton step must contract the residual quadratically
    s = JordanSolver(512, 64, comm=comm, device=dev, symm=mode)
    s.fill_formula(1)
    s.invert()
    X = s.local_inverse()
    X += 1e-9 * torch.randn_like(X)
    r0 = s.residual()
    s.refine(1)
    r1 = s.residual()
    if comm.rank == 0:
        results.append({"case": "refine", "r0": r0, "r1": r1, "ok": bool(r1 < 1e-2 * r0)})
    s.close()
    # tcgen05 update path across ranks
    s = JordanSolver(1024, 128, comm=comm, device=dev, symm=mode, precision="tc05")
    s.fill_formula(6)
    s.invert()
    X = s.gather_inverse()
    piv = s.pivots()
    if comm.rank == 0:
        A = formula_rows(6, 1024, torch.arange(1024))
        Xo, pivo = cpu_block_jordan(A, 128, p)
        results.append({"case": "tc05", "ok": bool(piv == pivo and float((X - Xo).abs().max() / Xo.abs().max()) < 1e-8)})
    s.close()
    # failure detection: rank 1 never joins the elimination -> the others must give up with
    # JB_ERR_TIMEOUT instead of hanging the GPU (watchdog in the cross-GPU spin-waits)
    os.environ["JB_SPIN_TIMEOUT_MS"] = "400"
    s = JordanSolver(256, 64, comm=comm, device=dev, symm=mode)
    os.environ.pop("JB_SPIN_TIMEOUT_MS")
    s.fill_formula(1)
    code = 0
    try:
        if comm.rank != 1:
            s.invert_async()
        s.status()          # collective verdict: every rank (also the idle one) learns about the timeout
    except _native.NativeError as exc:
        code = exc.code
    torch.cuda.synchronize()
    codes = comm.allgather_obj(code)
    # ... and the engine recovers: status() reset the flags on every rank, a normal inversion follows
    s.fill_formula(1)
    s.invert()
    X = s.gather_inverse()
    if comm.rank == 0:
        ref = torch.linalg.inv(formula_rows(1, 256, torch.arange(256)))
        results.append({"case": "watchdog+recover", "codes": codes,
                        "ok": bool(all(c == _native.JB_ERR_TIMEOUT for c in codes)
                                   and float((X - ref).abs().max()) < 1e-9 * float(ref.abs().max()))})
    s.close()
    if comm.rank == 0:
        print("MPRESULT " + json.dumps(results), flush=True)
    comm.barrier()


if __name__ == "__main__":
    main()
