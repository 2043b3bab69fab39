"""Pins the CPU oracle (oracle/oracle.c) against the known-answer vectors of SURVEY.md section 10
(K0 dlarnv, K2 b1_ss SpMV, K4 3x3-Laplacian ILU0/levels/solves, K5 size identities).  The reference
holds no golden vectors for SpMV/ILU0/SpTRSV (its arithmetic is in cuSPARSE/MKL/SuperLU): for those
"parity unpinned" beyond these KATs."""
import os

import numpy as np
import pytest

import orc
from respasol_b200 import host

GOLD = os.path.join(os.path.dirname(__file__), "golden")


def lap3():
    a = host.gen_laplace2d(3)
    assert a.rowptr.tolist() == [0, 3, 7, 10, 14, 19, 23, 26, 30, 33]
    assert a.colidx.tolist() == [0, 1, 3, 0, 1, 2, 4, 1, 2, 5, 0, 3, 4, 6, 1, 3, 4, 5, 7, 2, 4, 5, 8, 3, 6, 7, 4, 6, 7, 8, 5, 7, 8]
    return a


def test_k0_dlarnv():
    x = orc.dlarnv(6)
    assert x.tolist() == [0.12062469795087694, 0.6438459108216854, 0.06234171577016312, 0.49027924967339587,
                          0.3060786549148311, 0.816413585842529]
    assert np.array_equal(orc.dlarnv(1000), host.dlarnv_u01(1000))


def test_k2_b1ss_spmv():
    a = host.load_matrix_market(os.path.join(GOLD, "b1_ss.mtx"), 0, 0)
    assert a.rowptr.tolist() == [0, 3, 5, 7, 9, 11, 13, 15]
    assert a.colidx.tolist() == [1, 2, 3, 1, 4, 2, 5, 3, 6, 0, 4, 0, 5, 0, 6]
    y1 = orc.spmv(a, np.ones(7))
    np.testing.assert_allclose(y1, [3, -0.55, -0.9, -0.55, 0.96400058, 0.9823629, 0.992278221], rtol=0, atol=1e-15)
    x = orc.dlarnv(7)
    y = orc.spmv(a, x)
    ref = [1.1964668762652444, -0.50611051611001145, 0.019299642814089787, -0.041548033327566514,
           0.30173623575092434, 0.81428611598229961, 0.99624904350899124]
    assert y.tolist() == ref
    y32 = orc.spmv(a, x, "s")
    ref32 = np.array([1.19646692, -0.506110549, 0.019299645, -0.0415480733, 0.301736236, 0.814286113, 0.99624902], dtype=np.float32)
    assert np.array_equal(y32, ref32)
    ym = orc.spmv(a, x, "m")
    np.testing.assert_allclose(ym, ref, rtol=0, atol=2e-7)  # fp32-rounded A, fp64 accumulate
    # alpha/beta and the OpenMP baseline form agree with the sequential form
    y0 = np.arange(7, dtype=np.float64)
    yab = orc.spmv(a, x, alpha=2.0, beta=-0.5, y=y0)
    np.testing.assert_allclose(yab, 2.0 * y - 0.5 * y0, rtol=1e-15)
    assert np.array_equal(orc.spmv(a, x, threads=2), y)


def test_k2_b1ss_ilu0_structural_pivot():
    a = host.load_matrix_market(os.path.join(GOLD, "b1_ss.mtx"), 0, 0)
    st, row, _ = orc.ilu0(a)
    assert (st, row) == (5, 0)  # A(0,0) is missing


def test_k4_lap3_ilu0_levels_solves():
    a = lap3()
    st, row, lu = orc.ilu0(a)
    assert st == 0 and row == -1
    ref = [4, -1, -1, -0.25, 3.75, -1, -1, -0.26666666666666666, 3.7333333333333334, -1, -0.25, 3.75, -1, -1,
           -0.26666666666666666, -0.26666666666666666, 3.466666666666667, -1, -1, -0.26785714285714285,
           -0.28846153846153844, 3.443681318681319, -1, -0.26666666666666666, 3.7333333333333334, -1,
           -0.28846153846153844, -0.26785714285714285, 3.443681318681319, -1, -0.29038691663342636,
           -0.29038691663342636, 3.4192261667331474]
    assert lu.tolist() == ref
    levL, lpL, rowsL = orc.levels(a, 0)
    levU, lpU, rowsU = orc.levels(a, 1)
    assert levL.tolist() == [0, 1, 2, 1, 2, 3, 2, 3, 4]
    assert levU.tolist() == [4, 3, 2, 3, 2, 1, 2, 1, 0]
    assert lpL.tolist() == [0, 1, 3, 6, 8, 9] and rowsL.tolist() == [0, 1, 3, 2, 4, 6, 5, 7, 8]
    st, z, y = orc.sptrsv_lu(a, lu, np.ones(9))
    assert st == 0
    # SURVEY 10-K4 solved without FMA; the oracle (and the CUDA path) use explicit fma: <= 1 ulp apart
    np.testing.assert_allclose(z, [1, 1.25, 1.3333333333333333, 1.25, 1.6666666666666665, 1.8379120879120878,
                          1.3333333333333333, 1.8379120879120876, 2.067411248504188], rtol=2.3e-16, atol=0)
    np.testing.assert_allclose(y, [0.6082803881450638, 0.7165607762901277, 0.5471301913205786, 0.7165607762901275,
                          0.8899727197673999, 0.7092860475968267, 0.5471301913205786, 0.7092860475968267,
                          0.6046430237984133], rtol=4.5e-16, atol=0)


def test_k1_k3_identity_and_lower():
    one = host.load_matrix_market(os.path.join(GOLD, "one.mtx"), 0, 0)
    st, _, lu = orc.ilu0(one)
    assert st == 0 and np.array_equal(lu, one.values)
    assert orc.levels(one, 0)[0].tolist() == [0] * 7
    b = host.load_matrix_market(os.path.join(GOLD, "bcspwr01.mtx"), 0, 0)
    assert (b.m, b.nnz, b.stored) == (39, 131, 85)  # the reference's symmetric-file defect
    st, _, lu = orc.ilu0(b)
    assert st == 0 and np.array_equal(lu, b.values)  # lower triangular, unit diagonal: unchanged
    st, z, y = orc.sptrsv_lu(b, lu, np.ones(39))
    assert np.array_equal(y, z)                      # U = I


@pytest.mark.parametrize("N", [4, 17])
def test_k5_sizes_and_wavefronts(N):
    a = host.gen_laplace2d(N)
    assert a.stored == 5 * N * N - 4 * N
    lev = orc.levels(a, 0)[0].reshape(N, N)
    i, j = np.meshgrid(np.arange(N), np.arange(N), indexing="ij")
    assert np.array_equal(lev, i + j)
    s = host.gen_stencil27(5, 6, 7)
    assert s.stored == (3 * 5 - 2) * (3 * 6 - 2) * (3 * 7 - 2)


def test_zero_pivot_numeric():
    # 2x2 with u_11 = 1 - (1*1)/1 = 0
    a = host.CSR(2, 2, 4, np.array([0, 2, 4], np.int32), np.array([0, 1, 0, 1], np.int32), np.array([1.0, 1.0, 1.0, 1.0]))
    st, row, lu = orc.ilu0(a)
    assert (st, row) == (6, 1) and lu.tolist() == [1, 1, 1, 0]


def test_row_partition():
    a = host.gen_laplace2d(20)
    p = orc.row_partition(a, 4, True)
    assert p[0] == 0 and p[-1] == a.m and np.all(np.diff(p) > 0)
    for g in range(1, 4):
        t = g * a.stored // 4
        assert a.rowptr[p[g]] >= t and a.rowptr[p[g] - 1] < t
    assert orc.row_partition(a, 3, False).tolist() == [0, 134, 267, 400]


def test_residual_property_ilu0_exact_on_tridiagonal():
    # ILU0 of a tridiagonal matrix is its exact LU: L U y = b must reproduce b
    n = 50
    rp, ci, v = [0], [], []
    for i in range(n):
        for j in (i - 1, i, i + 1):
            if 0 <= j < n:
                ci.append(j); v.append(4.0 if i == j else -1.0 - 0.01 * i)
        rp.append(len(ci))
    a = host.CSR(n, n, len(ci), np.array(rp, np.int32), np.array(ci, np.int32), np.array(v))
    _, _, lu = orc.ilu0(a)
    b = orc.dlarnv(n)
    _, _, y = orc.sptrsv_lu(a, lu, b)
    np.testing.assert_allclose(orc.spmv(a, y), b, rtol=1e-13, atol=1e-14)


def test_oracle_pcg_solves_spd_systems():
    """orc_pcg_d (checker of rsg_pcg_d): converges to the scipy direct solution, ILU(0) cuts the iteration count,
    fp32 factors inside the fp64 iteration cost at most a couple of extra iterations."""
    import scipy.sparse as sp
    import scipy.sparse.linalg as spl
    from respasol_b200 import host
    g = host.gen_femlike(12)  # structurally symmetric, values not: take the symmetric part (still diagonally dominant)
    G = sp.csr_matrix((g.values, g.colidx, g.rowptr), shape=(g.m, g.n))
    M = ((G + G.T) * 0.5).tocsr()
    M.sort_indices()
    a = host.CSR(g.m, g.n, M.nnz, M.indptr.astype(np.int32), M.indices.astype(np.int32), M.data.copy(), 0)
    assert abs(M - M.T).max() < 1e-14
    b = orc.dlarnv(a.m) - 0.5
    x_direct = spl.spsolve(M.tocsc(), b)
    its = {}
    for pre in (None, np.float64, np.float32):
        lu = None if pre is None else orc.ilu0(a, pre)[2]
        x, it, rel = orc.pcg(a, b, lu, rtol=1e-11)
        assert rel <= 1e-11 and np.abs(x - x_direct).max() <= 1e-8 * np.abs(x_direct).max()
        its[pre] = it
    assert its[np.float64] < its[None] and its[np.float32] <= its[np.float64] + 2


def test_oracle_bicgstab_solves_unsymmetric_systems():
    import scipy.sparse as sp
    import scipy.sparse.linalg as spl
    from respasol_b200 import host
    a = host.gen_femlike(12)
    M = sp.csr_matrix((a.values, a.colidx, a.rowptr), shape=(a.m, a.n))
    assert abs(M - M.T).max() > 0.1  # unsymmetric values: CG does not apply
    b = orc.dlarnv(a.m) - 0.5
    x_direct = spl.spsolve(M.tocsc(), b)
    its = {}
    for pre in (None, np.float64, np.float32):
        lu = None if pre is None else orc.ilu0(a, pre)[2]
        x, it, rel = orc.bicgstab(a, b, lu, rtol=1e-11)
        assert rel <= 1e-11 and np.abs(x - x_direct).max() <= 1e-8 * np.abs(x_direct).max()
        its[pre] = it
    assert its[np.float64] < its[None] and its[np.float32] <= its[np.float64] + 3


def test_oracle_gmres_solves_unsymmetric_systems():
    """orc_gmres_d (restarted GMRES(30), right-preconditioned) against a sparse direct solve and scipy's own GMRES;
    enough iterations without a preconditioner to go through several restarts."""
    import scipy.sparse as sp
    import scipy.sparse.linalg as spl
    from respasol_b200 import host
    a = host.gen_femlike(12)
    M = sp.csr_matrix((a.values, a.colidx, a.rowptr), shape=(a.m, a.n))
    b = orc.dlarnv(a.m) - 0.5
    x_direct = spl.spsolve(M.tocsc(), b)
    its = {}
    for pre in (None, np.float64, np.float32):
        lu = None if pre is None else orc.ilu0(a, pre)[2]
        x, it, rel = orc.gmres(a, b, lu, rtol=1e-11)
        assert rel <= 1e-11 and np.abs(x - x_direct).max() <= 1e-8 * np.abs(x_direct).max(), (pre, it, rel)
        assert np.linalg.norm(b - M @ x) <= 1e-10 * np.linalg.norm(b)
        its[pre] = it
    assert its[np.float64] < its[None] and its[np.float32] <= its[np.float64] + 2
    # iteration cap: the estimate after exactly 7 Arnoldi steps, x updated with the partial cycle
    x7, it7, rel7 = orc.gmres(a, b, None, max_iters=7, rtol=1e-11)
    assert it7 == 7 and 0 < rel7 < 1
    assert abs(np.linalg.norm(b - M @ x7) / np.linalg.norm(b) - rel7) <= 1e-10
    # scipy's GMRES(30) without a preconditioner takes the same number of inner steps to the same tolerance (+-2)
    cnt = [0]
    spl.gmres(M, b, rtol=1e-11, atol=0.0, restart=30, maxiter=100, callback=lambda r: cnt.__setitem__(0, cnt[0] + 1),
              callback_type="pr_norm")
    assert abs(cnt[0] - its[None]) <= 2, (cnt[0], its[None])
