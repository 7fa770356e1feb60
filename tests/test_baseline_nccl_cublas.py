"""The in-tree NCCL + cuBLAS baseline (baseline/nccl_cublas/jordan_nccl.py) computes the same inverse
and the same pivot sequence as the CPU oracle - on CPU tensors here (single process and 2 ranks over
gloo); on the GPU box the same code runs on NCCL + cuBLAS/cuSOLVER."""
import json
import os
import subprocess
import sys

import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "baseline", "nccl_cublas"))

from mpi_jordan_crazy_acceleration_b200.models.eager import formula_rows  # noqa: E402
from mpi_jordan_crazy_acceleration_b200.ops import cpu_block_jordan  # noqa: E402


@pytest.mark.parametrize("fid,n,m", [(1, 100, 7), (6, 96, 12), (3, 24, 4), (5, 130, 32), (1, 30, 64)])
def test_single_process_matches_oracle(fid, n, m):
    from jordan_nccl import NcclCublasJordan

    b = NcclCublasJordan(n, m, device="cpu")
    b.fill_formula(fid)
    b.invert()
    A = formula_rows(fid, n, torch.arange(n))
    Xo, piv = cpu_block_jordan(A, min(m, n), 1)
    assert b.pivots == piv
    X = b.gather_inverse()
    assert (X - Xo).abs().max() <= 1e-9 * Xo.abs().max()


def test_singular_verdict():
    from jordan_nccl import NcclCublasJordan, SingularMatrix

    b = NcclCublasJordan(16, 4, device="cpu")
    b.W[:] = 1.0
    b.norm_a = 16.0
    with pytest.raises(SingularMatrix):
        b.invert()


WORKER = r'''
import json, os, sys
sys.path.insert(0, %(root)r); sys.path.insert(0, os.path.join(%(root)r, "baseline", "nccl_cublas"))
import torch, torch.distributed as dist
from jordan_nccl import NcclCublasJordan
dist.init_process_group("gloo")
out = []
for fid, n, m in [(1, 100, 7), (6, 96, 12), (6, 144, 8), (5, 130, 32)]:
    b = NcclCublasJordan(n, m, device="cpu")
    b.fill_formula(fid)
    b.invert()
    X = b.gather_inverse()
    if dist.get_rank() == 0:
        out.append({"fid": fid, "n": n, "m": m, "piv": b.pivots, "X": X.tolist()})
if dist.get_rank() == 0:
    print("RESULT " + json.dumps(out))
dist.barrier()
'''


@pytest.mark.parametrize("world", [2, 3])
def test_gloo_ranks_match_oracle(world, tmp_path):
    script = tmp_path / "w.py"
    script.write_text(WORKER % {"root": ROOT})
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}",
           "--master-addr", "127.0.0.1", "--master-port", str(29760 + world), str(script)]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=600, env=dict(os.environ, OMP_NUM_THREADS="1"))
    line = [ln for ln in r.stdout.splitlines() if ln.startswith("RESULT ")]
    assert line, r.stdout[-2000:] + r.stderr[-2000:]
    swaps = 0
    for rec in json.loads(line[-1][7:]):
        A = formula_rows(rec["fid"], rec["n"], torch.arange(rec["n"]))
        Xo, piv = cpu_block_jordan(A, rec["m"], world)
        assert rec["piv"] == piv
        swaps += sum(1 for t, g in enumerate(piv) if g != t and g % world != t % world)
        assert (torch.tensor(rec["X"], dtype=torch.float64) - Xo).abs().max() <= 1e-9 * Xo.abs().max()
    assert swaps > 0   # the cross-rank send/recv swap was exercised
