"""Multi-process CPU tests (gloo, world_size > 1) of the host-side distributed logic: block-cyclic
scatter of a text file (C4), corner gather (C5), min-loc tie rules across ranks (C9), pivot-row
broadcast (C10), cross-rank block-row swap (C11), error agreement (C3)."""
import os
import subprocess
import sys

import pytest
import torch

from mpi_jordan_crazy_acceleration_b200.models.eager import formula_rows
from mpi_jordan_crazy_acceleration_b200.ops import cpu_block_jordan
from mpi_jordan_crazy_acceleration_b200.utils import io as jio

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

WORKER = r'''
import json, os, sys
sys.path.insert(0, %(root)r)
import torch
from mpi_jordan_crazy_acceleration_b200.parallel.comm import Comm
from mpi_jordan_crazy_acceleration_b200.models.eager import EagerJordan
from mpi_jordan_crazy_acceleration_b200.parallel.layout import Layout
from mpi_jordan_crazy_acceleration_b200.utils import io as jio
comm = Comm.init_from_env(prefer_cuda=False)
out = []
for fid, n, m in [(1, 100, 7), (6, 96, 12), (3, 24, 4), (1, 30, 8)]:
    e = EagerJordan(n, m, comm=comm)
    e.fill_formula(fid)
    e.invert()
    X = e.gather_inverse()
    res = e.residual()
    if comm.rank == 0:
        out.append({"fid": fid, "n": n, "m": m, "piv": e.pivots, "X": X.tolist(), "res": res})
path = sys.argv[1]
lay = Layout(60, 8, comm.world, comm.rank)
rows = jio.scatter_text_matrix(path, lay, comm)
e = EagerJordan(60, 8, comm=comm)
e.upload_local_rows(rows)
e.invert()
X = e.gather_inverse()
c = jio.corner_from_rows(e.local_inverse(), lay, comm)
err = None
try:
    jio.scatter_text_matrix(path + ".missing", lay, comm)
except jio.MatrixFileError as exc:
    err = exc.code
codes = comm.allgather_obj(err)
# descriptor passing between the rank processes (carries CUDA VMM handles on the GPU box)
import tempfile
tf = tempfile.TemporaryFile()
tf.write(b"rank%%d" %% comm.rank); tf.flush()
fds = comm.exchange_fds(tf.fileno(), tag="t1")
seen = [os.pread(fd, 16, 0).decode() for fd in fds]
bfd = comm.bcast_fd(tf.fileno(), src=comm.world - 1, tag="t2")
seen.append(os.pread(bfd, 16, 0).decode())
fd_ok = comm.allgather_obj(seen == ["rank%%d" %% r for r in range(comm.world)] + ["rank%%d" %% (comm.world - 1)])
if comm.rank == 0:
    out.append({"fd_ok": fd_ok})
    out.append({"file": True, "X": X.tolist(), "corner": c.tolist(), "piv": e.pivots, "codes": codes})
    print("RESULT " + json.dumps(out))
'''


@pytest.mark.parametrize("world", [2, 3])
def test_eager_gloo_world(world, tmp_path):
    torch.manual_seed(11)
    A = torch.randn(60, 60, dtype=torch.float64)
    path = tmp_path / "m.txt"
    jio.write_matrix(str(path), A)
    script = tmp_path / "worker.py"
    script.write_text(WORKER % {"root": ROOT})
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}",
           "--master-addr", "127.0.0.1", "--master-port", str(29700 + world), str(script), str(path)]
    env = dict(os.environ, OMP_NUM_THREADS="1")
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=600, env=env)
    line = [ln for ln in r.stdout.splitlines() if ln.startswith("RESULT ")]
    assert line, r.stdout[-2000:] + r.stderr[-2000:]
    import json

    res = json.loads(line[-1][7:])
    assert res[-2] == {"fd_ok": [True] * world}
    for rec in res[:-2]:
        A_f = formula_rows(rec["fid"], rec["n"], torch.arange(rec["n"]))
        Xo, piv = cpu_block_jordan(A_f, rec["m"], world)      # oracle emulating `world` ranks
        assert rec["piv"] == piv
        assert torch.allclose(torch.tensor(rec["X"], dtype=torch.float64), Xo, atol=1e-9)
        assert rec["res"] < 1e-8
    f = res[-1]
    Xo, piv = cpu_block_jordan(A, 8, world)
    assert f["piv"] == piv
    assert torch.allclose(torch.tensor(f["X"], dtype=torch.float64), Xo, atol=1e-9)
    assert torch.allclose(torch.tensor(f["corner"], dtype=torch.float64), Xo[:10, :10], atol=1e-9)
    assert f["codes"] == [-1] * world          # every rank agreed on "cannot open"


def test_cli_under_torchrun_gloo_matches_single_process(tmp_path):
    """The CLI as an SPMD program (2 ranks, gloo): rank 0 prints the reference's report, file input is
    scattered block row by block row, and the numbers match the single-process run."""
    torch.manual_seed(3)
    n, m = 48, 6
    A = formula_rows(1, n, torch.arange(n)) + 0.25 * torch.eye(n, dtype=torch.float64)
    path = tmp_path / "a.txt"
    jio.write_matrix(str(path), A)
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2", "--master-addr",
           "127.0.0.1", "--master-port", "29741", "-m", "mpi_jordan_crazy_acceleration_b200.cli", str(n), str(m),
           str(path), "--backend", "eager"]
    env = dict(os.environ, OMP_NUM_THREADS="1", PYTHONPATH=ROOT)
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=600, env=env, cwd=ROOT)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    lines = [ln for ln in r.stdout.splitlines() if ln.strip() != "" or True]
    assert lines[0] == "A" and sum(ln.startswith("glob_time:") for ln in lines) == 1   # only rank 0 prints
    res = float([ln for ln in lines if ln.startswith("residual:")][0].split()[1])
    assert res < 1e-10
    import io as _io

    from mpi_jordan_crazy_acceleration_b200 import cli

    buf = _io.StringIO()
    assert cli.main([str(n), str(m), str(path), "--backend", "eager"], out=buf) == 0
    one = [ln for ln in buf.getvalue().splitlines() if "\t" in ln]
    two = [ln for ln in lines if "\t" in ln]
    assert len(one) == len(two) == 20
    a1 = torch.tensor([[float(v) for v in ln.split("\t") if v] for ln in one])
    a2 = torch.tensor([[float(v) for v in ln.split("\t") if v] for ln in two])
    assert torch.allclose(a1, a2, atol=0.011)
