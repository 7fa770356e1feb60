"""CLI / report / text-I/O contract vs the unmodified reference binary (built against the MPI shim)."""
import io
import os
import re
import subprocess
import sys

import pytest
import torch

from mpi_jordan_crazy_acceleration_b200 import cli
from mpi_jordan_crazy_acceleration_b200.models.eager import formula_rows
from mpi_jordan_crazy_acceleration_b200.utils import io as jio

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF = os.path.join(ROOT, "baseline", "_ref", "a.out")
MPIRUN = os.path.join(ROOT, "bin", "jb_mpirun")


@pytest.fixture(scope="module")
def reference():
    env = dict(os.environ, PYTHONPATH=ROOT)
    r = subprocess.run(["sh", os.path.join(ROOT, "baseline", "ref_cpu", "build_reference.sh")],
                       capture_output=True, text=True, env=env, cwd=ROOT)
    if r.returncode != 0 or not os.path.exists(REF):
        pytest.skip("reference source not available: " + r.stdout + r.stderr)
    return REF


def run_ref(np_, *args):
    r = subprocess.run([MPIRUN, "-np", str(np_), REF, *map(str, args)], capture_output=True, text=True, timeout=300)
    return r.returncode, r.stdout


def run_cli(*args):
    buf = io.StringIO()
    rc = cli.main([str(a) for a in args], out=buf)
    return rc, buf.getvalue()


def parse_report(text):
    lines = text.splitlines()
    assert lines[0] == "A"
    i = lines.index([ln for ln in lines if ln.startswith("glob_time:")][0])
    a = [[float(x) for x in ln.split("\t") if x] for ln in lines[1:i]]
    assert lines[i + 1] == "inverse matrix:" and lines[i + 2] == ""
    j = [k for k, ln in enumerate(lines) if ln.startswith("residual:")][0]
    x = [[float(v) for v in ln.split("\t") if v] for ln in lines[i + 3:j]]
    return a, x, float(lines[i].split()[1]), float(lines[j].split()[1])


def test_reference_config1_golden(reference):
    """BASELINE config 1: the reference as-is, mpirun -np 2, n=1200 m=60 -> residual 1.768067e-07."""
    rc, out = run_ref(2, 1200, 60)
    assert rc == 0
    a, x, t, res = parse_report(out)
    assert len(a) == 10 and len(x) == 10 and all(len(r) == 10 for r in a + x)
    assert re.search(r"^glob_time: \d+\.\d\d$", out, re.M)
    assert res == pytest.approx(1.768067e-07, rel=0.5)   # -Ofast bits are compiler dependent (D8)
    assert res <= 1.8e-07 * 1.5


def test_cli_report_matches_reference(reference):
    rc_ref, out_ref = run_ref(2, 1200, 60)
    rc, out = run_cli(1200, 60, "--backend", "eager")
    assert rc == rc_ref == 0
    a0, x0, _, res0 = parse_report(out_ref)
    a1, x1, _, res1 = parse_report(out)
    assert a0 == a1
    assert torch.allclose(torch.tensor(x0), torch.tensor(x1), atol=0.011)   # "%.2f" rounding
    assert res1 <= res0                                                     # accuracy bar (Appendix A.11)
    # line structure is identical
    strip = lambda s: [re.sub(r"[-0-9.e+]+", "#", ln) for ln in s.splitlines()]  # noqa: E731
    assert strip(out) == strip(out_ref)


def test_usage_and_error_paths(reference, tmp_path):
    rc_ref, out_ref = run_ref(2)
    rc, out = run_cli()
    assert rc == rc_ref == 1 and out.startswith("usage:") and out.rstrip().endswith("n m [<file>]")
    assert out_ref.startswith("usage:")
    assert run_cli(0, 5)[0] == 1 and run_cli(5, "x")[0] == 1 and run_cli(1, 2, 3, 4, 5, 6)[0] == 1
    rc_ref, out_ref = run_ref(2, 8, 4, "/nonexistent/file")
    rc, out = run_cli(8, 4, "/nonexistent/file", "--backend", "eager")
    assert (rc, out) == (2, "cannot open /nonexistent/file\n") and rc_ref == 2 and out_ref == out
    short = tmp_path / "short.txt"
    short.write_text("1 2 3 4 5 6 7")
    rc_ref, out_ref = run_ref(1, 3, 3, short)
    rc, out = run_cli(3, 3, short, "--backend", "eager")
    assert rc == rc_ref == 2 and out == out_ref == f"cannot read {short}\n"
    garbled = tmp_path / "garbled.txt"
    garbled.write_text("1 2 x 4")
    assert run_cli(2, 2, garbled, "--backend", "eager") == (2, f"cannot read {garbled}\n")
    ones = tmp_path / "ones.txt"
    ones.write_text(" ".join(["1"] * 16))
    rc_ref, out_ref = run_ref(2, 4, 2, ones)
    rc, out = run_cli(4, 2, ones, "--backend", "eager")
    assert rc == rc_ref == 2 and out.endswith("singular matrix\n") and out_ref.endswith("singular matrix\n")
    # Hilbert (formula 2): the reference build with -DHILBERT reports singular for n >= 16
    assert run_cli(64, 32, 10, 2, "--backend", "eager")[0] == 2
    r = subprocess.run([MPIRUN, "-np", "2", os.path.join(ROOT, "baseline", "_ref", "a_hilbert.out"), "64", "32"],
                       capture_output=True, text=True)
    assert r.returncode == 2 and r.stdout.endswith("singular matrix\n")


def test_file_input_same_as_reference(reference, tmp_path):
    torch.manual_seed(5)
    n, m = 60, 8
    A = formula_rows(1, n, torch.arange(n)) + torch.eye(n, dtype=torch.float64) * 0.5
    path = tmp_path / "a.txt"
    jio.write_matrix(str(path), A)
    assert torch.equal(jio.read_full(str(path), n), A)
    rc_ref, out_ref = run_ref(3, n, m, path)
    rc, out = run_cli(n, m, path, "--backend", "eager")
    assert rc == rc_ref == 0
    a0, x0, _, res0 = parse_report(out_ref)
    a1, x1, _, res1 = parse_report(out)
    assert a0 == a1 and torch.allclose(torch.tensor(x0), torch.tensor(x1), atol=0.011)
    assert res1 <= max(res0, 1e-9)


def test_extended_selectors():
    rc, out = run_cli(24, 4, 5, 3, "--backend", "eager")       # print size 5, anti-diagonal formula
    assert rc == 0
    a, x, _, res = parse_report(out)
    assert len(a) == 5 and len(a[0]) == 5 and res < 1e-12
    rc, out2 = run_cli(24, 4, "--print", 5, "--formula", 3, "--backend", "eager")
    strip = lambda s: re.sub(r"glob_time: .*", "", s)  # noqa: E731
    assert rc == 0 and strip(out2) == strip(out)


def test_format_corner_is_printf_exact():
    t = torch.tensor([[0.005, -0.004, 1234.567], [1e-9, -1e-9, 2.0]], dtype=torch.float64)
    assert jio.format_corner(t) == "0.01\t-0.00\t1234.57\t\n0.00\t-0.00\t2.00\t\n".replace("0.01", "%.2f" % 0.005)


def test_reference_bench_arm_line(reference):
    """bench.py --impl reference: unmodified reference through its own CLI, one JSON line."""
    import json

    env = dict(os.environ, JB_REF_BUDGET_S="30")
    r = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--gpus", "2",
                        "--n", "600", "--m", "60"], capture_output=True, text=True, timeout=300, env=env, cwd=ROOT)
    assert r.returncode == 0, r.stderr
    d = json.loads(r.stdout.strip().splitlines()[-1])
    assert d["impl"] == "reference" and d["metric"] == "fp64_inverse_gflops" and d["n_gpus"] == 2
    assert d["config"]["n"] == 600 and d["value"] > 0 and d["gpu_launches"] == 0
    # end to end for the reference = job launch -> glob_time line; it can only be slower than Jordan alone
    assert d["e2e"]["unit"] == "GFLOP/s" and 0 < d["e2e"]["value"] <= d["value"] * 1.01
    assert d["e2e"]["h2d_bytes_per_step"] == 0 and d["e2e"]["d2h_bytes_per_step"] == 0


def test_reference_bench_arm_quotes_best_block_size(reference):
    """m = 64 is off the reference's 3x3-blocked fast path: the arm also reports m = 63."""
    import json

    env = dict(os.environ, JB_REF_BUDGET_S="30")
    r = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--gpus", "2",
                        "--n", "640", "--m", "64"], capture_output=True, text=True, timeout=300, env=env, cwd=ROOT)
    assert r.returncode == 0, r.stderr
    d = json.loads(r.stdout.strip().splitlines()[-1])
    best = d["reference_best_block_size"]
    assert best["m"] == 63 and best["n"] % 63 == 0 and best["value"] > 0
