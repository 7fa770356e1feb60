import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from mpi_jordan_crazy_acceleration_b200.models.jordan import JordanSolver, _SingleComm
from mpi_jordan_crazy_acceleration_b200.models.eager import formula_rows
n, m = int(sys.argv[1]), int(sys.argv[2])
reps = int(sys.argv[3]) if len(sys.argv) > 3 else 3
s = JordanSolver(n, m, comm=_SingleComm())
ref = torch.linalg.inv(formula_rows(1, n, torch.arange(n)).cuda())
for it in range(reps):
    s.fill_formula(1)
    try:
        sec = s.invert()
        X = s.local_inverse()
        print(it, "ok sec", sec, "maxerr", float((X - ref).abs().max()), "piv ok", s.pivots() == list(range(s.layout.Nr)), flush=True)
    except Exception as e:
        print(it, "FAIL", e, "pivots", s.pivots()[:80], flush=True)
