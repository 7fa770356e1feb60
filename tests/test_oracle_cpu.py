"""CPU oracle + eager (in-place) algorithm: numerics, pivot rule, tie rules, singular verdicts."""
import pytest
import torch

from mpi_jordan_crazy_acceleration_b200 import _native
from mpi_jordan_crazy_acceleration_b200.models.eager import EagerJordan, formula_rows, minloc_fold
from mpi_jordan_crazy_acceleration_b200.ops import cpu_block_jordan, cpu_inverse_block, cpu_minloc_fold, cpu_residual


def gen(fid, n):
    return formula_rows(fid, n, torch.arange(n))


@pytest.mark.parametrize("n,m,p", [(120, 30, 1), (120, 7, 3), (100, 16, 2), (9, 20, 3), (30, 8, 8)])
def test_oracle_absdiff(n, m, p):
    A = gen(1, n)
    X, piv = cpu_block_jordan(A, m, p)
    assert piv == list(range(len(piv)))  # |i-j|: the diagonal block always wins (SURVEY D1 note)
    assert torch.allclose(X, torch.linalg.inv(A), rtol=0, atol=1e-9)
    assert cpu_residual(A, X) < 1e-8


@pytest.mark.parametrize("fid,n,m,p", [(6, 128, 8, 2), (3, 12, 3, 2), (3, 24, 4, 3), (6, 96, 12, 4)])
def test_pivoting_inputs_oracle_vs_inplace(fid, n, m, p):
    A = gen(fid, n)
    X, piv = cpu_block_jordan(A, m, p)
    assert any(pg != t for t, pg in enumerate(piv))  # real block-row swaps happen
    assert cpu_residual(A, X) < 1e-9
    e = EagerJordan(n, m)
    e.fill_formula(fid)
    e.invert()
    assert torch.allclose(e.local_inverse(), X, rtol=0, atol=1e-9)


def test_random_matrix_matches_torch():
    torch.manual_seed(0)
    A = torch.randn(60, 60, dtype=torch.float64)
    X, piv = cpu_block_jordan(A, 7, 3)
    assert torch.allclose(X, torch.linalg.inv(A), rtol=0, atol=1e-9)
    e = EagerJordan(60, 7)
    e.upload_local_rows(A)
    e.invert()
    assert e.pivots == piv or torch.allclose(e.local_inverse(), X, atol=1e-9)
    assert e.residual() < 1e-10


def test_singular_verdicts():
    with pytest.raises(_native.SingularMatrixError):
        cpu_block_jordan(torch.ones(4, 4, dtype=torch.float64), 2, 1)
    # Hilbert: works only for tiny n, singular by the 1e-15*||A|| test for n >= 16
    with pytest.raises(_native.SingularMatrixError):
        cpu_block_jordan(gen(2, 32), 32, 1)
    X, _ = cpu_block_jordan(gen(2, 8), 4, 1)
    assert cpu_residual(gen(2, 8), X) < 1e-2
    e = EagerJordan(4, 2)
    e.upload_local_rows(torch.ones(4, 4, dtype=torch.float64))
    with pytest.raises(_native.SingularMatrixError):
        e.invert()


def test_inverse_block_semantics():
    torch.manual_seed(1)
    a = torch.randn(17, 17, dtype=torch.float64)
    inv, sing = cpu_inverse_block(a, float(a.abs().sum(1).max()))
    assert not sing and torch.allclose(inv, torch.linalg.inv(a), atol=1e-10)
    z = torch.zeros(3, 3, dtype=torch.float64)
    assert cpu_inverse_block(z, 0.0)[1]          # |norm| < EPS
    tiny = torch.eye(3, dtype=torch.float64) * 1e-3
    assert cpu_inverse_block(tiny, 1e13)[1]      # |a_kk| < 1e-15 * norm
    assert not cpu_inverse_block(tiny, 1.0)[1]


def test_minloc_tie_rules():
    # ties across ranks -> HIGHEST tied rank (MPI rank-ordered fold of a non-commutative op)
    assert cpu_minloc_fold([1.0, 1.0, 1.0], [1, 1, 1]) == 2
    assert cpu_minloc_fold([1.0, 0.5, 0.5, 2.0], [1, 1, 1, 1]) == 2
    assert cpu_minloc_fold([0.5, 1.0, 1.0], [1, 1, 1]) == 0
    # non-singular beats singular regardless of the value
    assert cpu_minloc_fold([0.0, 3.0, 0.0], [0, 1, 0]) == 1
    assert cpu_minloc_fold([0.0, 0.0], [0, 0]) == -1
    for norms, ok in [([1.0, 1.0, 1.0], [1, 1, 1]), ([3.0, 2.0, 2.0, 5.0], [1, 1, 1, 0]), ([0.0, 0.0], [0, 0])]:
        assert minloc_fold([(a, b, 0) for a, b in zip(norms, ok)]) == cpu_minloc_fold(norms, ok)
