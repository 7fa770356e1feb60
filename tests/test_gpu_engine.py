"""GPU tests of the fused sm_100a engine against plain PyTorch fp64 references and the CPU oracle."""
import pytest
import torch

from mpi_jordan_crazy_acceleration_b200 import _native
from mpi_jordan_crazy_acceleration_b200.models.eager import formula_rows
from mpi_jordan_crazy_acceleration_b200.models.jordan import JordanSolver, _SingleComm, invert
from mpi_jordan_crazy_acceleration_b200.ops import block_inverse, cpu_block_jordan, cpu_inverse_block

pytestmark = pytest.mark.gpu


def gen(fid, n):
    return formula_rows(fid, n, torch.arange(n))


def solver(n, m, **kw):
    return JordanSolver(n, m, comm=_SingleComm(), **kw)


@pytest.mark.parametrize("m", [8, 32, 60, 128, 160, 256])
def test_block_inverse_matches_oracle(m):
    torch.manual_seed(m)
    k = 5
    blocks = torch.randn(k, m, m, dtype=torch.float64)
    blocks[1] = gen(1, m)                 # structured
    blocks[2, 3] = blocks[2, 4]           # exactly singular (two equal rows)
    blocks[3] = 0.0                       # zero block
    norm_a = float(blocks.abs().sum(dim=2).max())
    inv, norms, ok = block_inverse(blocks.cuda(), norm_a)
    assert _native.cuda_loaded()
    inv, norms, ok = inv.cpu(), norms.cpu(), ok.cpu()
    for b in range(k):
        ref, sing = cpu_inverse_block(blocks[b], norm_a)
        assert bool(ok[b]) == (not sing), f"block {b}: verdict differs"
        if not sing:
            scale = ref.abs().max()
            assert (inv[b] - ref).abs().max() <= 1e-9 * scale
            assert abs(norms[b] - ref.abs().sum(1).max()) <= 1e-9 * ref.abs().sum(1).max()


@pytest.mark.parametrize("n,m", [(512, 64), (1200, 60), (1000, 128), (300, 7), (768, 256), (200, 200), (96, 32)])
def test_invert_absdiff_vs_torch(n, m):
    s = solver(n, m)
    s.fill_formula("absdiff")
    s.invert()
    X = s.local_inverse().cpu()
    A = gen(1, n)
    ref = torch.linalg.inv(A)
    assert s.pivots() == list(range(s.layout.Nr))
    assert (X - ref).abs().max() <= 1e-9 * ref.abs().max()
    res = s.residual()
    assert res <= 1e-6, res
    host_res = float((A @ X - torch.eye(n, dtype=torch.float64)).abs().sum(1).max())
    assert abs(res - host_res) <= 1e-3 * max(host_res, 1e-12) + 1e-12
    s.close()


@pytest.mark.parametrize("fid,n,m", [(6, 512, 64), (3, 96, 8), (6, 1024, 128), (5, 500, 60), (4, 256, 32)])
def test_pivoting_inputs_vs_oracle(fid, n, m):
    A = gen(fid, n)
    Xo, piv = cpu_block_jordan(A, m, 1)
    s = solver(n, m)
    s.fill_formula(fid)
    s.invert()
    X = s.local_inverse().cpu()
    assert s.pivots() == piv
    assert (X - Xo).abs().max() <= 1e-8 * Xo.abs().max()
    assert s.residual() <= 1e-6 * max(1.0, float(A.abs().sum(1).max()) * float(Xo.abs().sum(1).max()) * 1e-4)
    s.close()


def test_uploaded_random_matrix_and_convenience():
    torch.manual_seed(3)
    n = 384
    A = torch.randn(n, n, dtype=torch.float64)
    X = invert(A, m=64).cpu()
    ref = torch.linalg.inv(A)
    assert (X - ref).abs().max() <= 1e-8 * ref.abs().max()
    Xo, piv = cpu_block_jordan(A, 64, 1)
    s = solver(n, 64, keep_copy=True)
    s.upload_local_rows(A.pin_memory())
    s.invert()
    assert s.pivots() == piv
    assert s.residual() <= 1e-9
    s.close()


def test_singular_and_recovery():
    s = solver(64, 16)
    s.upload_local_rows(torch.ones(64, 64, dtype=torch.float64))
    with pytest.raises(_native.SingularMatrixError):
        s.invert()
    # the same engine is usable afterwards
    s.fill_formula(1)
    s.invert()
    ref = torch.linalg.inv(gen(1, 64))
    assert (s.local_inverse().cpu() - ref).abs().max() <= 1e-10
    s.close()
    h = solver(256, 32)
    h.fill_formula("hilbert")  # reference verdict: singular matrix, rc 2
    with pytest.raises(_native.SingularMatrixError):
        h.invert()
    h.close()


def test_repeated_inversions_same_engine():
    s = solver(640, 128)
    ref = torch.linalg.inv(gen(1, 640))
    for _ in range(3):
        s.fill_formula(1)
        s.invert()
        assert (s.local_inverse().cpu() - ref).abs().max() <= 1e-9 * ref.abs().max()
    assert s.launch_count > 0
    s.close()


def test_refinement_reduces_residual():
    n, m = 512, 64
    s = solver(n, m)
    s.fill_formula(1)
    s.invert()
    X = s.local_inverse()
    X += 1e-9 * torch.randn_like(X)       # damage the inverse
    r0 = s.residual()
    s.refine(1)
    r1 = s.residual()
    assert r1 < r0 * 1e-2, (r0, r1)   # Newton-Schulz: quadratic contraction
    s.close()


@pytest.mark.parametrize("fid,n,m", [(1, 1024, 128), (6, 1024, 128), (1, 768, 256), (5, 640, 128)])
def test_tc05_update_path_matches_fp64_path(fid, n, m):
    """K3b: the tcgen05 int8-split trailing update gives fp64-level results (same pivots, same X)."""
    A = gen(fid, n)
    Xo, piv = cpu_block_jordan(A, m, 1)
    s = solver(n, m, precision="tc05")
    s.fill_formula(fid)
    s.invert()
    X = s.local_inverse().cpu()
    assert s.pivots() == piv
    assert (X - Xo).abs().max() <= 1e-8 * Xo.abs().max()
    r_tc = s.residual()
    s.close()
    f = solver(n, m)
    f.fill_formula(fid)
    f.invert()
    r_64 = f.residual()
    f.close()
    assert r_tc <= 10 * r_64 + 1e-12, (r_tc, r_64)


def test_tc05_gemm_building_block():
    from mpi_jordan_crazy_acceleration_b200.ops import tc05_gemm_update

    torch.manual_seed(0)
    L = torch.randn(256, 128, dtype=torch.float64, device="cuda")
    R = torch.randn(128, 192, dtype=torch.float64, device="cuda")
    C0 = torch.randn(256, 192, dtype=torch.float64, device="cuda")
    ref = C0 + L @ R
    out = tc05_gemm_update(C0.clone(), L, R, 8)
    assert ((out - ref).abs() / (L.abs() @ R.abs())).max() < 1e-13


def _invert_w(n, m, fid, env=None):
    import os

    old = {k: os.environ.get(k) for k in (env or {})}
    os.environ.update(env or {})
    try:
        s = solver(n, m)          # scheduling switches are read when the engine is created
    finally:
        for k, v in old.items():
            os.environ.pop(k, None) if v is None else os.environ.__setitem__(k, v)
    s.fill_formula(fid)
    s.invert()
    out = (s.local_inverse().clone(), s.pivots(), s.residual())
    s.close()
    return out


@pytest.mark.parametrize("fid,n,m", [(6, 9600, 32), (6, 4096, 128), (1, 2304, 256)])
def test_lookahead_schedule_is_bitwise_equal_to_serial(fid, n, m):
    """The look-ahead schedule (A / B1 / B2 with the next pivot row scaled and broadcast inside B2,
    K1(t+1) on its own stream) only reorders independent tiles, so it must reproduce the serial
    schedule bit for bit - also with more pivot candidates than the GPU has SMs (n = 9600, m = 32:
    300 candidate CTAs), where K1(t+1) runs in several waves beside the update and any read of a
    buffer K1 overwrites (the inverse of the chosen pivot block, the L panel) would show up here."""
    Xs, ps, rs = _invert_w(n, m, fid, {"JB_NO_LOOKAHEAD": "1"})
    # JB_SCHED: 1 = phase A in line, 2 = phase A beside B2 of the previous step; JB_B1_FRAC: share of
    # the tiles updated while K1 runs (0 and 1 are the degenerate splits)
    for env in ({"JB_SCHED": "1"}, {"JB_SCHED": "2"}, {"JB_SCHED": "2", "JB_B1_FRAC": "0.0"},
                {"JB_SCHED": "1", "JB_B1_FRAC": "1.0"}, {"JB_SCHED": "2", "JB_B1_FRAC": "1.0"},
                {"JB_SCHED": "2", "JB_K1_GATED": "1"}):
        Xl, pl, rl = _invert_w(n, m, fid, env)
        assert pl == ps, env
        assert torch.equal(Xl, Xs), (env, float((Xl - Xs).abs().max()))
        assert abs(rl - rs) <= 1e-9 * rs   # row sums are accumulated with atomics: last-bit noise
    if fid == 6:
        assert any(g != t for t, g in enumerate(ps))   # block-row swaps did happen
