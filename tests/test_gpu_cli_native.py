"""The native C++ driver bin/jordan (csrc/cli/main.cpp; reference main/solve, main.cpp:65-93,
343-519) launched through the in-tree MPI launcher, one rank per GPU: report contract, exit codes,
file input, agreement with the Python front end and with the unmodified reference binary, and the
native symmetric heap (VMM + descriptors over the shim's sockets + NVLS multicast) on >= 2 GPUs."""
import os
import re
import subprocess

import pytest
import torch

from mpi_jordan_crazy_acceleration_b200.models.eager import formula_rows
from mpi_jordan_crazy_acceleration_b200.utils import io as jio

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
MPIRUN = os.path.join(ROOT, "bin", "jb_mpirun")
JORDAN = os.path.join(ROOT, "bin", "jordan")
REF = os.path.join(ROOT, "baseline", "_ref", "a.out")


def ngpu():
    return torch.cuda.device_count()


def run_native(np_, *args, env=None, timeout=600):
    e = dict(os.environ)
    e.update(env or {})
    r = subprocess.run([MPIRUN, "-np", str(np_), JORDAN, *map(str, args)], capture_output=True, text=True,
                       timeout=timeout, env=e)
    return r.returncode, r.stdout, r.stderr


def parse_report(text):
    lines = text.splitlines()
    i0 = lines.index("A")
    i = [k for k, ln in enumerate(lines) if ln.startswith("glob_time:")][0]
    a = [[float(x) for x in ln.split("\t") if x] for ln in lines[i0 + 1:i]]
    assert lines[i + 1] == "inverse matrix:" and lines[i + 2] == ""
    j = [k for k, ln in enumerate(lines) if ln.startswith("residual:")][0]
    x = [[float(v) for v in ln.split("\t") if v] for ln in lines[i + 3:j]]
    return a, x, float(lines[i].split()[1]), float(lines[j].split()[1])


def rank_counts():
    return [p for p in (1, 2, 4, 8) if p <= max(1, ngpu())]


@pytest.mark.parametrize("p", [1, 2, 4, 8])
def test_report_contract_and_reference_agreement(p):
    """n = 1200, m = 60 (BASELINE config 1 shape): same report lines as the reference, same corner,
    residual at the reference's level (1.768e-07)."""
    if p > ngpu():
        pytest.skip(f"needs {p} GPUs")
    rc, out, err = run_native(p, 1200, 60)
    assert rc == 0, out + err
    a, x, t, res = parse_report(out)
    assert len(a) == 10 and len(x) == 10 and all(len(r) == 10 for r in a + x)
    assert re.search(r"^glob_time: \d+\.\d\d$", out, re.M)
    assert res <= 1.8e-07 * 1.5
    A = formula_rows(1, 1200, torch.arange(1200))
    X = torch.linalg.inv(A)
    want_a = [[round(float(v), 2) for v in row[:10]] for row in A[:10]]
    want_x = [[round(float(v), 2) + 0.0 for v in row[:10]] for row in X[:10]]
    assert a == want_a
    for r1, r2 in zip(x, want_x):
        for v1, v2 in zip(r1, r2):
            assert abs(v1 - v2) <= 0.011
    if os.path.exists(REF) and p >= 2:   # the reference skips the residual for p == 1 (main.cpp:499-513)
        r = subprocess.run([MPIRUN, "-np", str(p), REF, "1200", "60"], capture_output=True, text=True, timeout=600)
        ra, rx, _, rres = parse_report(r.stdout)
        assert ra == a
        for r1, r2 in zip(rx, x):
            for v1, v2 in zip(r1, r2):
                assert abs(v1 - v2) <= 0.011
        assert res <= rres * 1.5


def test_usage_and_error_exit_codes(tmp_path):
    rc, out, _ = run_native(1)
    assert rc == 1 and out.startswith("usage:")                     # main.cpp:77-83
    rc, out, _ = run_native(1, 0, 4)
    assert rc == 1 and out.startswith("usage:")
    rc, out, _ = run_native(1, 8, 4, tmp_path / "missing.txt")
    assert rc == 2 and "cannot open" in out                          # main.cpp:392
    short = tmp_path / "short.txt"
    short.write_text("1 2 3\n4 5\n")
    rc, out, _ = run_native(1, 3, 2, short)
    assert rc == 2 and "cannot read" in out                          # main.cpp:394
    ones = tmp_path / "ones.txt"
    jio.write_matrix(str(ones), torch.ones(4, 4, dtype=torch.float64))
    rc, out, _ = run_native(1, 4, 2, ones)
    assert rc == 2 and "singular matrix" in out                      # main.cpp:437-438
    rc, out, _ = run_native(1, 256, 32, 10, 2)                       # Hilbert (config 5): same verdict
    assert rc == 2 and "singular matrix" in out


def test_block_size_larger_than_matrix_and_unsupported_block_size():
    """The reference accepts m > n (one block row); a block size the kernels cannot hold is reported as
    such, not as an allocation failure."""
    rc, out, err = run_native(1, 96, 500)
    assert rc == 0, out + err
    assert parse_report(out)[3] < 1e-9
    rc, out, _ = run_native(1, 2000, 1000)
    assert rc == 2 and "unsupported configuration" in out and "Not enough memory" not in out


@pytest.mark.parametrize("p", [1, 2, 8])
def test_file_input_matches_python_front_end(p, tmp_path):
    if p > ngpu():
        pytest.skip(f"needs {p} GPUs")
    torch.manual_seed(11)
    n, m = 150, 16
    A = torch.randn(n, n, dtype=torch.float64) + 4.0 * torch.eye(n, dtype=torch.float64)
    path = tmp_path / "a.txt"
    jio.write_matrix(str(path), A)
    rc, out, err = run_native(p, n, m, path)
    assert rc == 0, out + err
    a, x, _, res = parse_report(out)
    X = torch.linalg.inv(A)
    for got, want in ((a, A), (x, X)):
        for i in range(10):
            for j in range(10):
                assert abs(got[i][j] - float(want[i, j])) <= 0.006
    assert res < 1e-9


@pytest.mark.parametrize("p", [2, 4, 8])
def test_native_heap_with_multicast_and_pivoting(p):
    """>= 2 GPUs: the symmetric region comes from the native VMM heap; with NVSwitch the kernels store
    through the multicast alias.  Formula 6 forces cross-GPU block-row swaps; both symmetric-memory
    modes must give the same report."""
    if p > ngpu():
        pytest.skip(f"needs {p} GPUs")
    outs = {}
    for mode in ("vmm", "ipc"):
        rc, out, err = run_native(p, 18 * 64, 64, 10, 6, env={"JB_SYMM": mode, "JB_VERBOSE": "1"})
        assert rc == 0, out + err
        outs[mode] = out
        assert parse_report(out)[3] < 1e-6
    assert "symmetric region: vmm" in outs["vmm"]
    assert "symmetric region: cuda-ipc" in outs["ipc"]
    strip = lambda s: [ln for ln in s.splitlines() if not ln.startswith(("symmetric", "glob_time", "seconds"))]  # noqa: E731
    assert strip(outs["vmm"]) == strip(outs["ipc"])
