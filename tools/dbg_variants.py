"""torchrun debug worker: residual of one inversion under several schedule switches, in ONE process
group.  usage: torchrun ... tools/dbg_variants.py n m [fid]"""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from mpi_jordan_crazy_acceleration_b200.models.jordan import JordanSolver
from mpi_jordan_crazy_acceleration_b200.parallel.comm import Comm
comm = Comm.init_from_env()
dev = int(os.environ.get("LOCAL_RANK", "0"))
torch.cuda.set_device(dev)
n, m = int(sys.argv[1]), int(sys.argv[2])
fid = int(sys.argv[3]) if len(sys.argv) > 3 else 1
KEYS = ("JB_SCHED", "JB_K1_GATED", "JB_B2_OVERLAP", "JB_NO_LOOKAHEAD", "JB_B1_FRAC", "JB_NO_MULTICAST")
VARIANTS = [{}, {"JB_SCHED": "1"}, {"JB_K1_GATED": "1"}, {"JB_B2_OVERLAP": "1"}, {"JB_NO_LOOKAHEAD": "1"},
            {"JB_B1_FRAC": "0.06"}, {"JB_NO_MULTICAST": "1"}, {}]
for env in VARIANTS:
    for k in KEYS:
        os.environ.pop(k, None)
    os.environ.update(env)
    s = JordanSolver(n, m, comm=comm, device=dev)
    s.fill_formula(fid)
    sec = s.invert()
    res = s.residual()
    if comm.rank == 0:
        print("DBGV", json.dumps({"env": env, "sec": round(sec, 4), "residual": res, "symm": s.symm_mode}), flush=True)
    s.close()
