"""ILU0 / level sets / SpTRSV parity through the C ABI against the CPU oracle.

Integers (diagonal positions via zero-pivot reports, level_ptr, level_rows) must be bit-exact.
Factors and thread-per-row solves follow the oracle's operation order with explicit fma and IEEE
division, so they are compared BIT-EXACT as well; the stated fallback tolerance is the
componentwise bound c*eps*(|L||U|)_ij, which bit-exactness trivially meets."""
import os

import numpy as np
import pytest
import torch

import orc
from respasol_b200 import host
from respasol_b200.gpu import DeviceCSR, ZERO_PIVOT_NUMERIC, ZERO_PIVOT_STRUCTURAL

pytestmark = pytest.mark.gpu
GOLD = os.path.join(os.path.dirname(__file__), "golden")


def tridiag(n):
    rp, ci, v = [0], [], []
    for i in range(n):
        for j in (i - 1, i, i + 1):
            if 0 <= j < n:
                ci.append(j); v.append(4.0 if i == j else -1.0 - 0.01 * (i % 7))
        rp.append(len(ci))
    return host.CSR(n, n, len(ci), np.array(rp, np.int32), np.array(ci, np.int32), np.array(v))


def arrow(n):
    """dense last row and column + diagonal: one very long row (beyond the staged-row capacity)"""
    rp, ci, v = [0], [], []
    for i in range(n - 1):
        ci += [i, n - 1]; v += [4.0 + i % 3, -1.0 / (1 + i % 5)]
        rp.append(len(ci))
    ci += list(range(n)); v += [-0.5 / (1 + i % 4) for i in range(n - 1)] + [float(n)]
    rp.append(len(ci))
    return host.CSR(n, n, len(ci), np.array(rp, np.int32), np.array(ci, np.int32), np.array(v))


def to_base1(a):
    return host.CSR(a.m, a.n, a.nnz, a.rowptr + 1, a.colidx + 1, a.values.copy(), 1)


MATS = {
    "one": lambda: host.load_matrix_market(os.path.join(GOLD, "one.mtx")),
    "bcspwr01": lambda: host.load_matrix_market(os.path.join(GOLD, "bcspwr01.mtx")),
    "lap2d_3": lambda: host.gen_laplace2d(3),
    "lap2d_40": lambda: host.gen_laplace2d(40),
    "lap2d_40_base1": lambda: to_base1(host.gen_laplace2d(40)),
    "femlike_10": lambda: host.gen_femlike(10),
    "femlike_20": lambda: host.gen_femlike(20),
    "stencil27_12": lambda: host.gen_stencil27(12, seed=2),
    "tridiag_3000": lambda: tridiag(3000),
    "arrow_700": lambda: arrow(700),
    "powerlaw_5k": lambda: host.gen_powerlaw(5000, 2.2, 48, 2, 900, 4),
}


@pytest.mark.parametrize("name", sorted(MATS))
def test_levels_bit_exact(handle, name):
    a = MATS[name]()
    A = DeviceCSR(a, "d")
    handle.ilu0_analysis(A)
    for which in (0, 1):
        lp, rows = handle.levels(which, a.m)
        _, lp_ref, rows_ref = orc.levels(a, which)
        assert np.array_equal(lp, lp_ref)
        assert np.array_equal(rows, rows_ref)


@pytest.mark.parametrize("name", sorted(MATS))
@pytest.mark.parametrize("dtype", [np.float64, np.float32])
@pytest.mark.parametrize("trsv_algo", [0, 1, 2, 3])
def test_ilu0_and_solves(handle, name, dtype, trsv_algo):
    a = MATS[name]()
    prec = "d" if dtype == np.float64 else "s"
    A = DeviceCSR(a, prec)
    st_ref, piv_ref, lu_ref = orc.ilu0(a, dtype)
    st, piv = handle.ilu0_factor(A)
    assert (st, piv) == (st_ref, piv_ref)
    lu = A.values.cpu().numpy()
    assert np.array_equal(lu.view(np.uint8), lu_ref.view(np.uint8)), "ILU0 factors are not bit-identical to the oracle"
    b = (orc.dlarnv(a.m) + 0.5).astype(dtype)
    _, z_ref, y_ref = orc.sptrsv_lu(a, lu_ref, b)
    handle.set_option("TRSV_ALGO", trsv_algo)
    try:
        z, y = handle.sptrsv_lu(A, A.values, torch.from_numpy(b).cuda())
    finally:
        handle.set_option("TRSV_ALGO", 0)
    assert np.array_equal(z.cpu().numpy().view(np.uint8), z_ref.view(np.uint8))
    assert np.array_equal(y.cpu().numpy().view(np.uint8), y_ref.view(np.uint8))


def test_golden_lap3(handle):
    a = host.gen_laplace2d(3)
    A = DeviceCSR(a, "d")
    handle.ilu0_factor(A)
    ref = [4, -1, -1, -0.25, 3.75, -1, -1, -0.26666666666666666, 3.7333333333333334, -1, -0.25, 3.75, -1, -1,
           -0.26666666666666666, -0.26666666666666666, 3.466666666666667, -1, -1, -0.26785714285714285,
           -0.28846153846153844, 3.443681318681319, -1, -0.26666666666666666, 3.7333333333333334, -1,
           -0.28846153846153844, -0.26785714285714285, 3.443681318681319, -1, -0.29038691663342636,
           -0.29038691663342636, 3.4192261667331474]
    assert A.values.cpu().tolist() == ref
    lp, rows = handle.levels(0, 9)
    assert lp.tolist() == [0, 1, 3, 6, 8, 9] and rows.tolist() == [0, 1, 3, 2, 4, 6, 5, 7, 8]
    z, y = handle.sptrsv_lu(A, A.values, torch.ones(9, dtype=torch.float64, device="cuda"))
    np.testing.assert_allclose(y.cpu().numpy(), [0.6082803881450638, 0.7165607762901277, 0.5471301913205786,
                                                 0.7165607762901275, 0.8899727197673999, 0.7092860475968267,
                                                 0.5471301913205786, 0.7092860475968267, 0.6046430237984133], rtol=4.5e-16)


def test_structural_zero_pivot_b1ss(handle):
    a = host.load_matrix_market(os.path.join(GOLD, "b1_ss.mtx"))
    A = DeviceCSR(a, "d")
    before = A.values.clone()
    st, row = handle.ilu0_analysis(A)
    assert (st, row) == (ZERO_PIVOT_STRUCTURAL, 0)       # harness prints "A(0,0) is missing"
    st, row = handle.ilu0_factor(A)
    assert (st, row) == (ZERO_PIVOT_STRUCTURAL, 0) and torch.equal(before, A.values)
    # level sets are still defined and must match the oracle
    lp, rows = handle.levels(0, a.m)
    _, lp_ref, rows_ref = orc.levels(a, 0)
    assert np.array_equal(lp, lp_ref) and np.array_equal(rows, rows_ref)


def test_numeric_zero_pivot(handle):
    a = host.CSR(2, 2, 4, np.array([0, 2, 4], np.int32), np.array([0, 1, 0, 1], np.int32), np.array([1.0, 1.0, 1.0, 1.0]))
    A = DeviceCSR(a, "d")
    st, row = handle.ilu0_factor(A)
    assert (st, row) == (ZERO_PIVOT_NUMERIC, 1)          # harness prints "L(1,1) is zero"
    assert A.values.cpu().tolist() == [1, 1, 1, 0]


def test_wavefront_rule_lap2d(handle):
    N = 100
    a = host.gen_laplace2d(N)
    A = DeviceCSR(a, "d")
    handle.ilu0_analysis(A)
    lp, rows = handle.levels(0, a.m)
    assert len(lp) - 1 == 2 * N - 1
    lev = np.empty(a.m, dtype=np.int64)
    for l in range(len(lp) - 1):
        lev[rows[lp[l]:lp[l + 1]]] = l
    i, j = np.meshgrid(np.arange(N), np.arange(N), indexing="ij")
    assert np.array_equal(lev.reshape(N, N), i + j)


def test_host_pointer_entry_points(handle):
    a = host.gen_femlike(9)
    st, piv, lu = handle.ilu0_factor_host(a)
    _, _, lu_ref = orc.ilu0(a)
    assert st == 0 and np.array_equal(lu, lu_ref)
    b = orc.dlarnv(a.m)
    y = handle.sptrsv_lu_host(a, lu, b)
    assert np.array_equal(y, orc.sptrsv_lu(a, lu_ref, b)[2])


def test_repeated_factorizations_reuse_the_plan(handle):
    a = host.gen_femlike(12)
    A = DeviceCSR(a, "d")
    _, _, lu_ref = orc.ilu0(a)
    pristine = A.values.clone()
    for _ in range(3):
        A.values.copy_(pristine)
        st, _ = handle.ilu0_factor(A)
        assert st == 0 and np.array_equal(A.values.cpu().numpy(), lu_ref)


def test_config3_full_size(handle):
    """BASELINE config 3: FEM-like 47^3 = 103 823 rows, ~1.6 M nnz (2cubes_sphere-shaped)."""
    a = host.gen_femlike(47, 0.58, 3)
    assert a.m == 103823 and 1.4e6 < a.stored < 1.8e6
    for dtype, prec in ((np.float64, "d"), (np.float32, "s")):
        A = DeviceCSR(a, prec)
        st, piv = handle.ilu0_factor(A)
        st_ref, _, lu_ref = orc.ilu0(a, dtype)
        assert st == st_ref == 0
        assert np.array_equal(A.values.cpu().numpy(), lu_ref)
        b = np.ones(a.m, dtype=dtype)
        z, y = handle.sptrsv_lu(A, A.values, torch.from_numpy(b).cuda())
        _, z_ref, y_ref = orc.sptrsv_lu(a, lu_ref, b)
        assert np.array_equal(z.cpu().numpy(), z_ref) and np.array_equal(y.cpu().numpy(), y_ref)
        for which in (0, 1):
            lp, rows = handle.levels(which, a.m)
            _, lp_ref, rows_ref = orc.levels(a, which)
            assert np.array_equal(lp, lp_ref) and np.array_equal(rows, rows_ref)
    # size-independent properties.  (1) the defining one of ILU(0): (L*U)_ij = a_ij on the pattern of A.
    import scipy.sparse as sp
    A = DeviceCSR(a, "d")
    lu = A.values.clone()
    handle.ilu0_factor(A, lu)
    luv = lu.cpu().numpy()
    rp, ci = a.rowptr.astype(np.int64), a.colidx.astype(np.int64)
    rows = np.repeat(np.arange(a.m), np.diff(rp))
    low = ci < rows
    L = sp.csr_matrix((np.where(low, luv, 0.0), ci, rp), shape=(a.m, a.m)) + sp.identity(a.m, format="csr")
    U = sp.csr_matrix((np.where(low, 0.0, luv), ci, rp), shape=(a.m, a.m))
    P = (L @ U).tocsr()
    on_pattern = np.asarray(P[rows, ci]).ravel()
    absP = (abs(L) @ abs(U)).tocsr()
    bound = 40 * 2.0 ** -53 * np.asarray(absP[rows, ci]).ravel()
    assert np.all(np.abs(on_pattern - a.values) <= bound)
    # (2) L U y = b is solved to rounding: r = b - L(U y) against |L||U||y|
    x = torch.from_numpy(orc.dlarnv(a.m)).cuda()
    Ax = handle.spmv(A, x)
    _, y = handle.sptrsv_lu(A, lu, Ax)
    yh = y.cpu().numpy()
    r = Ax.cpu().numpy() - L @ (U @ yh)
    assert np.all(np.abs(r) <= 200 * 2.0 ** -53 * (abs(L) @ (abs(U) @ np.abs(yh))) + 1e-300)
    # (3) M^-1 A is a contraction towards the identity for this diagonally dominant matrix
    assert float((y - x).abs().max() / x.abs().max()) < 1.0


def test_hung_solve_reports_dependency_timeout(handle):
    """A right-hand side that carries the pending sentinel (a NaN payload no arithmetic produces) makes row 0 look
    unwritten for ever: its dependants give up after RSG_OPT_WAIT_SPINS sweeps and the next synchronising call
    returns RSG_DEPENDENCY_TIMEOUT instead of RSG_OK (round 1 returned RSG_OK with NaNs in y)."""
    from respasol_b200.gpu import RsgError
    a = host.gen_laplace2d(20)
    A = DeviceCSR(a, "d")
    handle.ilu0_factor(A)
    b = torch.ones(a.m, dtype=torch.float64, device="cuda")
    sentinel = torch.tensor([0xFFF8DEADBEEF0001 - (1 << 64)], dtype=torch.int64, device="cuda").view(torch.float64)
    b[0:1] = sentinel
    handle.set_option("WAIT_SPINS", 2000)
    handle.set_option("TRSV_ALGO", 1)  # the flag-polling kernel (the chain kernel has no flags: it would just propagate the NaN)
    try:
        handle.sptrsv_lu(A, A.values, b)
        with pytest.raises(RsgError) as ei:
            handle.synchronize()
        assert ei.value.status == 7  # RSG_DEPENDENCY_TIMEOUT
        handle.synchronize()  # reported once
        good = torch.ones(a.m, dtype=torch.float64, device="cuda")
        z, y = handle.sptrsv_lu(A, A.values, good)
        handle.synchronize()
        assert bool(torch.isfinite(y).all())
    finally:
        handle.set_option("WAIT_SPINS", 0)
        handle.set_option("TRSV_ALGO", 0)


def test_unsorted_rows_are_refused(handle):
    """ILU0 / SpTRSV split a row at its diagonal and binary-search it: a row whose columns do not ascend is refused with
    RSG_BAD_ARG by the analysis, factor and solve calls (the header states the contract); SpMV takes any order."""
    from respasol_b200.gpu import RsgError
    a = host.gen_laplace2d(12)
    ci = a.colidx.copy()
    vals = a.values.copy()
    r = 40
    k0, k1 = int(a.rowptr[r]), int(a.rowptr[r + 1])
    ci[k0:k1] = ci[k0:k1][::-1]
    vals[k0:k1] = vals[k0:k1][::-1]
    b = host.CSR(a.m, a.n, a.nnz, a.rowptr, ci, vals, a.base)
    B = DeviceCSR(b, "d")
    for call in (lambda: handle.ilu0_analysis(B), lambda: handle.sptrsv_analysis(B), lambda: handle.ilu0_factor(B),
                 lambda: handle.sptrsv_lu(B, B.values, torch.ones(b.m, dtype=torch.float64, device="cuda"))):
        with pytest.raises(RsgError) as ei:
            call()
        assert ei.value.status == 1  # RSG_BAD_ARG
    x = orc.dlarnv(a.n)
    y = handle.spmv(B, torch.from_numpy(x).cuda()).cpu().numpy()
    assert np.abs(y - orc.spmv(a, x, "d")).max() <= 1e-12
